"""Pattern, mirroring src/pattern.rs:21-45 (Pattern::new / matches / original / is_case_insensitive). The regex engine
is the library's own DFA compiler (Rust-`regex` syntax subset); the same DFA is what the device runs."""
import ctypes as C

from . import _cabi


class Pattern:
    def __init__(self, pattern, case_insensitive=False):
        if pattern == "":
            raise ValueError("Pattern cannot be empty")                      # src/pattern.rs:22-24
        err = C.create_string_buffer(256)
        rc = _cabi.lib().vgen_pattern_check(pattern.encode(), int(case_insensitive), err, 256)
        if rc != 0:
            raise ValueError("Invalid regex pattern: %s: %s" % (pattern, err.value.decode()))   # src/pattern.rs:32-33
        self._pattern = pattern
        self._ci = bool(case_insensitive)

    @staticmethod
    def new(pattern, case_insensitive=False):
        return Pattern(pattern, case_insensitive)

    def matches(self, address):                                              # src/pattern.rs:43-45
        return _cabi.lib().vgen_pattern_is_match(self._pattern.encode(), int(self._ci), address.encode()) == 1

    def original(self):
        return self._pattern

    def is_case_insensitive(self):
        return self._ci

    def validate_charset(self, fmt):                                          # src/pattern.rs:48-175
        buf = C.create_string_buffer(512)
        n = _cabi.lib().vgen_pattern_validate_charset(int(fmt), self._pattern.encode(), int(self._ci), buf, 512)
        if n < 0:
            raise _cabi.VgenError(n, "vgen_pattern_validate_charset")
        return list(buf.value.decode())

    def estimate_difficulty(self, fmt):                                       # src/pattern.rs:177-253
        return int(_cabi.lib().vgen_pattern_estimate_difficulty(int(fmt), self._pattern.encode(), int(self._ci)))

    def lowering(self, fmt):
        """How the scan path will test this pattern on the device for `fmt`: dict(kind=..., intervals=[(lo, hi), ...], exact=..., suffix=(mask, value) or None, confirm="dfa" or None, residue=(58^k, value, ...) or None)."""
        buf = C.create_string_buffer(4096)
        rc = _cabi.lib().vgen_pattern_lowering(int(fmt), self._pattern.encode(), int(self._ci), buf, 4096)
        if rc < 0:
            raise _cabi.VgenError(rc, "vgen_pattern_lowering")
        parts = buf.value.decode().split(";")
        out = {"kind": parts[0], "intervals": [], "exact": None, "suffix": None, "confirm": None, "residue": None}
        for p in parts[1:]:
            if p.startswith("exact="):
                out["exact"] = bytes.fromhex(p[6:])
            elif p.startswith("confirm="):      # "dfa": the device runs the DFA on the keys that pass the interval prefilter
                out["confirm"] = p[8:]
            elif p.startswith("residue="):      # (modulus 58^k, value, ...): Base58 literal suffix, tested on the payload mod 58^k
                out["residue"] = tuple(int(x, 16) for x in p[8:].split(":"))
            elif p.startswith("suffix="):       # (mask, value) over the Bech32 checksum bits / the last 32 digest bits
                out["suffix"] = tuple(int(x, 16) for x in p[7:].split(":"))
            elif p:
                lo, hi = p.split("-")
                out["intervals"].append((int(lo, 16), int(hi, 16)))
        return out

    def device_dfa(self, fmt):
        """The dense DFA the device runs for `fmt`: dict(table=bytes (uint16 LE), accept=bytes, sym_class=bytes, n_states, n_classes, start, flags)."""
        L = _cabi.lib()
        info = (C.c_uint32 * 4)()
        sym = C.create_string_buffer(64)
        rc = L.vgen_pattern_dfa(int(fmt), self._pattern.encode(), int(self._ci), None, 0, None, 0, sym, info)
        if rc not in (0, -5):
            raise _cabi.VgenError(rc, "vgen_pattern_dfa")
        ns, nc = info[0], info[1]
        table = C.create_string_buffer(2 * ns * nc)
        accept = C.create_string_buffer(ns)
        rc = L.vgen_pattern_dfa(int(fmt), self._pattern.encode(), int(self._ci), table, ns * nc, accept, ns, sym, info)
        if rc != 0:
            raise _cabi.VgenError(rc, "vgen_pattern_dfa")
        return dict(table=table.raw, accept=accept.raw, sym_class=sym.raw, n_states=ns, n_classes=nc, start=info[2], flags=info[3])
