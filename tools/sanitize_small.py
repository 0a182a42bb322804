"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck): every format's digest path, the three matcher
kinds, the degenerate-batch path (range touching scalar 0) and the P2TR stages, on reference-sized windows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
import vgen_b200 as vg

r = vg.GpuRunner(1 << 18)
for fmt in ("p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "p2tr", "ethereum"):
    n = 3000
    got = r.digest_batch(fmt, 1, n)
    raw = oracle.digest_batch(fmt, 1, n)
    st = 32 if fmt == "p2tr" else 20
    assert got == b"".join(raw[32 * i:32 * i + st] for i in range(n)), fmt
for fmt, pat in (("p2pkh", "^1A"), ("p2pkh", "ab$"), ("p2wpkh", "qq$"), ("ethereum", "^0xAb"), ("p2tr", "^bc1pq"), ("p2tr", "zz$"), ("p2sh-p2wpkh", "^3A")):
    start, n = 0, (1 << 18) + 999
    res = vg.scan_gpu_with_runner(vg.Pattern(pat), vg.ScanConfig(format=vg.AddressFormat.parse(fmt), count=0, start=start, end=start + n - 1), None, None, r)
    keys, ops, _ = oracle.scan_range(fmt, pat, start, n)
    assert [int(m.hex, 16) for m in res.matches] == keys and res.operations == ops, (fmt, pat)
addr = oracle.address("p2pkh", 70000)[0]
res = vg.scan_gpu_with_runner(vg.Pattern("^" + addr + "$"), vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=0, start=1, end=1 << 18), None, None, r)
assert [int(m.hex, 16) for m in res.matches] == [70000]
r.close()
print("sanitize_small ok")
