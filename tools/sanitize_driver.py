#!/usr/bin/env python
"""Workload for compute-sanitizer (tools/sanitize.sh): short scans that drive every kernel of the library through every matcher
mode — k_walk / k_slow for the six formats in interval, exact and DFA mode (incl. the Base58 / Bech32 / EIP-55 device encoders
and the literal-suffix prefilters), k_slow through ranges that contain scalar 0 and n-1, k_taptweak with a range end inside a
warp, the table builders, k_points_of_scalars, and the dump kernels. One CTA per window keeps the run short under the tools.
Prints one line per scan; exits non-zero on any library error. Needs no oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import vgen_b200 as vg  # noqa: E402

N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
S = 1 << 64
PREFIX = {"p2pkh": "^1C", "p2pkh-uncompressed": "^1C", "p2wpkh": "^bc1qa", "p2sh-p2wpkh": "^3C", "p2tr": "^bc1pq", "ethereum": "^0xab"}
SUFFIX = {"p2pkh": "at$", "p2pkh-uncompressed": "[0-9]{2}$", "p2wpkh": "qq$", "p2sh-p2wpkh": "Zz$", "p2tr": "a.*7$", "ethereum": "Ab$"}
MIXED = {"p2pkh": "^1C.*z$", "p2wpkh": "^bc1qa.*zz$", "ethereum": "^0xAb"}


def scan(r, fmt, pat, start, n, count=vg.UNLIMITED):
    cfg = vg.ScanConfig(format=vg.AddressFormat.parse(fmt), count=count, start=start, end=start + n - 1)
    res = vg.scan_gpu_with_runner(vg.Pattern(pat), cfg, None, None, r)
    print("%-20s %-28s %-10s keys %8d hits %6d" % (fmt, pat[:28], r.matcher_kind(), res.operations, len(res.matches)), flush=True)
    return res


def main():
    full = "--full" in sys.argv                   # default geometry: 296 CTAs, 77.6 M keys per window
    n = 1 << 18
    r = vg.GpuRunner(0 if full else 262144)       # else one 256-thread CTA per window
    if full:
        for fmt in PREFIX:
            scan(r, fmt, PREFIX[fmt] + ("at" if fmt.startswith("p2pkh") else ""), S, r.batch_size + 12345)
            if fmt != "p2tr":
                scan(r, fmt, SUFFIX[fmt] if fmt != "p2pkh-uncompressed" else "99$", S, r.batch_size // 4)
        st = r.stats()
        print("stats", st, flush=True)
        assert st["damaged_guards"] == 0, st
        r.close()
        print("SANITIZE_DRIVER_OK")
        return
    for fmt in PREFIX:
        scan(r, fmt, PREFIX[fmt], S, n)
        scan(r, fmt, SUFFIX[fmt], S, n)
        if fmt in MIXED:
            scan(r, fmt, MIXED[fmt], S, n)
        # exact: the address of a key inside the range, taken from a dense scan of its neighbourhood
        one = scan(r, fmt, ".", S + 1000, 4, count=1)
        res = scan(r, fmt, "^" + one.matches[0].address.replace(".", "\\.") + "$", S, 2000)
        assert [int(m.hex, 16) for m in res.matches] == [S + 1000], res.matches
    # degenerate batches (k_slow): the range contains scalar 0 / reaches n - 1
    scan(r, "p2pkh", "^1", 0, 3000)
    scan(r, "ethereum", "^0x", N - 3000, 3000)
    scan(r, "p2tr", "^bc1p", 0, 3000)
    # P2TR range end inside a warp of the tweak stage, and inside a CTA
    scan(r, "p2tr", "q$", S, 64 * 37 + 5)
    scan(r, "p2tr", "q$", S, 64 * 256 * 3 + 64 * 40 + 1)
    # dump kernels (generate_batch / digest_batch / internal keys), ragged sizes
    assert len(r.generate_batch(12345, 4097)) == 4097
    for fmt in PREFIX:
        assert len(r.digest_batch(fmt, S + 7, 1025)) == 1025 * (32 if fmt == "p2tr" else 20)
    r.dispatch_p2tr(S, 0, n=777)
    assert len(r.await_result_p2tr(0)[0]) == 777
    static_batch(r)
    st = r.stats()
    print("stats", st, flush=True)
    assert st["damaged_guards"] == 0, st
    r.close()
    print("SANITIZE_DRIVER_OK")


def static_batch(r):
    """the allocation-free entry point (static_batch.cu): same kernels, work area in a __device__ array; output into torch memory"""
    import ctypes as C
    import torch
    from vgen_b200 import _cabi
    L = _cabi.lib()
    err = C.create_string_buffer(256)
    for start, n in [(0, 5000), (N - 3000, 5000), (S + 99, 1 << 20), (S, 1)]:
        out = torch.full((n * 20,), 0xEE, dtype=torch.uint8, device="cuda:0")
        assert L.vgen_cuda_hash160_batch_static(0, start.to_bytes(32, "big"), n, out.data_ptr(), None, err, len(err)) == 0, err.value
        got = out.cpu().numpy().tobytes()
        assert got == b"".join(r.generate_batch(start, n)), (hex(start), n)
        viol = (C.c_uint32 * 2)(0, 0)
        if hasattr(L, "vgen_cuda_static_violations"):           # checked build only
            assert L.vgen_cuda_static_violations(0, viol) == 0 and list(viol) == [0, 0], list(viol)
        print("%-20s %-28s %-10s keys %8d violations %s" % ("static hash160", hex(start)[:28], "dump", n, list(viol)), flush=True)


if __name__ == "__main__":
    main()
