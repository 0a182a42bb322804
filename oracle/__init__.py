"""ORACLE — ctypes loader for the CPU restatement (oracle/vgen_oracle.c). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (vgen_b200) never does.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libvgen_oracle.so")

FORMATS = {"p2pkh": 0, "p2pkh-uncompressed": 1, "p2wpkh": 2, "p2sh-p2wpkh": 3, "p2tr": 4, "ethereum": 5}


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("vgen_oracle.c", "fe.h", "hashes.h", "encode.h", "regex.h")]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        u8p = C.POINTER(C.c_uint8)
        L.vgo_pubkey.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
        L.vgo_address.argtypes = [C.c_int, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p]
        L.vgo_address_from_digest.argtypes = [C.c_int, C.c_char_p, C.c_char_p]
        L.vgo_hash160_batch.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p]
        L.vgo_digest_batch.argtypes = [C.c_int, C.c_char_p, C.c_uint64, C.c_void_p]
        for f in (L.vgo_sha256, L.vgo_ripemd160, L.vgo_keccak256):
            f.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p]
            f.restype = None
        L.vgo_base58check.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        L.vgo_base58_decode.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.vgo_segwit.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p]
        L.vgo_base58.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        L.vgo_taproot_output.argtypes = [C.c_char_p, C.c_char_p]
        L.vgo_eip55.argtypes = [C.c_char_p, C.c_char_p]
        L.vgo_eip55.restype = None
        L.vgo_regex_check.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        L.vgo_regex_match.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        L.vgo_scan_range.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_uint64, C.c_int,
                                     C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                     C.POINTER(C.c_double)]
        L.vgo_init()
        _lib = L
    return _lib


def _fmt(f):
    return FORMATS[f] if isinstance(f, str) else int(f)


def key_bytes(k):
    return k if isinstance(k, (bytes, bytearray)) else int(k).to_bytes(32, "big")


def pubkey(key):
    x = C.create_string_buffer(32); y = C.create_string_buffer(32)
    if lib().vgo_pubkey(key_bytes(key), x, y):
        return None
    return int.from_bytes(x.raw, "big"), int.from_bytes(y.raw, "big")


def address(fmt, key):
    """AddressGenerator::generate (address.rs:92-151) -> (address, wif, digest) or None."""
    a = C.create_string_buffer(128); w = C.create_string_buffer(128); d = C.create_string_buffer(32)
    if lib().vgo_address(_fmt(fmt), key_bytes(key), a, w, d):
        return None
    n = 32 if _fmt(fmt) == 4 else 20
    return a.value.decode(), w.value.decode(), d.raw[:n]


def address_from_digest(fmt, digest):
    a = C.create_string_buffer(128)
    lib().vgo_address_from_digest(_fmt(fmt), bytes(digest).ljust(32, b"\0"), a)
    return a.value.decode()


def hash160_batch(start, n):
    buf = C.create_string_buffer(20 * n)
    lib().vgo_hash160_batch(key_bytes(start), n, buf)
    return buf.raw


def digest_batch(fmt, start, n):
    buf = C.create_string_buffer(32 * n)
    lib().vgo_digest_batch(_fmt(fmt), key_bytes(start), n, buf)
    return buf.raw


def sha256(m):
    o = C.create_string_buffer(32); lib().vgo_sha256(m, len(m), o); return o.raw


def ripemd160(m):
    o = C.create_string_buffer(20); lib().vgo_ripemd160(m, len(m), o); return o.raw


def keccak256(m):
    o = C.create_string_buffer(32); lib().vgo_keccak256(m, len(m), o); return o.raw


def base58check(m):
    o = C.create_string_buffer(256); lib().vgo_base58check(m, len(m), o); return o.value.decode()


def base58(m):
    o = C.create_string_buffer(2 * len(m) + 8); lib().vgo_base58(m, len(m), o); return o.value.decode()


def base58_decode(s, cap=128):
    o = C.create_string_buffer(cap); n = lib().vgo_base58_decode(s.encode(), o, cap); return None if n < 0 else o.raw[:n]


def taproot_output(xonly):
    """BIP341 key-path output key (no script tree) of an x-only internal key, or None if x is not on the curve."""
    o = C.create_string_buffer(32)
    return None if lib().vgo_taproot_output(bytes(xonly), o) else o.raw


def segwit(witver, prog):
    o = C.create_string_buffer(128); lib().vgo_segwit(witver, prog, len(prog), o); return o.value.decode()


def eip55(addr20):
    o = C.create_string_buffer(64); lib().vgo_eip55(addr20, o); return o.value.decode()


def regex_check(pattern, ci=False):
    """None if the pattern compiles, else the error string."""
    err = C.create_string_buffer(128)
    rc = lib().vgo_regex_check(pattern.encode(), int(ci), err, 128)
    return None if rc == 0 else err.value.decode()


def regex_match(pattern, s, ci=False):
    r = lib().vgo_regex_match(pattern.encode(), int(ci), s.encode())
    if r < 0:
        raise ValueError("invalid pattern: %r" % pattern)
    return bool(r)


def scan_range(fmt, pattern, start, n_keys, ci=False, threads=None, cap=1 << 20):
    """scan_range_cpu (scanner.rs:211-330) over [start, start+n_keys-1].
    Returns (sorted list of matching secret ints, ops, elapsed_secs)."""
    threads = threads or os.cpu_count() or 1
    out = C.create_string_buffer(32 * cap)
    found = C.c_uint64(0); ops = C.c_uint64(0); el = C.c_double(0)
    rc = lib().vgo_scan_range(_fmt(fmt), pattern.encode(), int(ci), key_bytes(start), n_keys, threads,
                              out, cap, C.byref(found), C.byref(ops), C.byref(el))
    if rc:
        raise ValueError("invalid pattern: %r" % pattern)
    if found.value > cap:
        raise OverflowError("more matches than cap")
    raw = out.raw[:32 * found.value]
    keys = sorted(int.from_bytes(raw[32 * i:32 * i + 32], "big") for i in range(found.value))
    return keys, ops.value, el.value
