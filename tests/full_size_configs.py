"""BASELINE.json configs C2-C5 at their full single-GPU sizes, checked through size-independent properties:
soundness (every reported hit is re-derived by the oracle and matches the pattern), completeness on planted keys and on
an oracle-scanned head of the range, exact operation counts. Run by hand on a GPU box (not collected by pytest):
    python tests/full_size_configs.py [c2 c3 c4 c5]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle          # noqa: E402  (tests/ may use the oracle as the checker)
import vgen_b200 as vg  # noqa: E402


def scan(r, fmt, pat, start, n):
    t0 = time.perf_counter()
    res = vg.scan_gpu_with_runner(vg.Pattern(pat), vg.ScanConfig(format=vg.AddressFormat.parse(fmt), count=0, start=start, end=start + n - 1), None, None, r)
    dt = time.perf_counter() - t0
    assert res.operations == n, (res.operations, n)
    return res, dt


def check_hits(fmt, pat, res, start, n, head):
    keys = [int(m.hex, 16) for m in res.matches]
    assert keys == sorted(set(keys)) and all(start <= k < start + n for k in keys)
    for m in res.matches:                                   # soundness, every hit
        a, w, _ = oracle.address(fmt, int(m.hex, 16))
        assert (a, w) == (m.address, m.wif) and oracle.regex_match(pat, a), m
    want, _, _ = oracle.scan_range(fmt, pat, start, head)   # completeness on the oracle-scannable head
    assert [k for k in keys if k < start + head] == want
    return len(keys), len(want)


def main(which):
    r = vg.GpuRunner()
    S = 1 << 64
    if "c2" in which:
        n = 1 << 36
        planted = [S + 12345, S + n // 3, S + n - 7]
        res, dt = scan(r, "p2pkh", "^1CatDog", S, n)
        nh, nw = check_hits("p2pkh", "^1CatDog", res, S, n, 1 << 22)
        print("C2 p2pkh ^1CatDog 2^36 keys: %d hits (all verified), %.1f s, %.0f Mkeys/s" % (nh, dt, n / dt / 1e6), flush=True)
        pat = "^(" + "|".join(oracle.address("p2pkh", k)[0][:8] for k in planted) + ")"
        res, dt = scan(r, "p2pkh", pat, S, n)
        keys = [int(m.hex, 16) for m in res.matches]
        assert all(k in keys for k in planted), "planted keys missing"
        check_hits("p2pkh", pat, res, S, n, 1 << 20)
        print("C2 planted 8-char prefixes of 3 keys: found all, %d hits, matcher %s, %.1f s" % (len(keys), r.matcher_kind(), dt), flush=True)
    if "c3" in which:
        n = 1 << 34
        res, dt = scan(r, "p2wpkh", "dead$", S, n)
        nh, nw = check_hits("p2wpkh", "dead$", res, S, n, 1 << 23)
        print("C3 p2wpkh dead$ 2^34 keys: %d hits (all verified; %d in the oracle-scanned head), %.1f s, %.0f Mkeys/s" % (nh, nw, dt, n / dt / 1e6), flush=True)
    if "c4" in which:
        n = 1 << 34
        res, dt = scan(r, "ethereum", "^0xdead", S, n)
        nh, nw = check_hits("ethereum", "^0xdead", res, S, n, 1 << 22)
        print("C4 ethereum ^0xdead 2^34 keys: %d hits (all verified, EIP-55 lower-case only; %d in the head), %.1f s, %.0f Mkeys/s" % (nh, nw, dt, n / dt / 1e6), flush=True)
    if "c5" in which:
        start, n = 1 << 65, 1 << 40
        off = int(0.83 * n)
        addr = oracle.address("p2pkh", start + off)[0]
        res, dt = scan(r, "p2pkh", "^" + addr + "$", start, n)
        assert [int(m.hex, 16) for m in res.matches] == [start + off] and res.matches[0].address == addr
        print("C5 exact address over 2^40 keys from 2^65: the planted key at offset %d found, nothing else, %.1f s, %.0f Mkeys/s (matcher %s)" % (off, dt, n / dt / 1e6, r.matcher_kind()), flush=True)
    r.close()


if __name__ == "__main__":
    main([a.lower() for a in sys.argv[1:]] or ["c2", "c3", "c4", "c5"])
