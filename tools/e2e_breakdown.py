"""Times the phases of scan_gpu_with_runner on the bench workload (tools only; not part of the product path)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgen_b200 as vg

pat_s = sys.argv[1] if len(sys.argv) > 1 else "^1Cat"
fmt = vg.AddressFormat.parse(sys.argv[2] if len(sys.argv) > 2 else "p2pkh")
wps = int(sys.argv[3]) if len(sys.argv) > 3 else 32
r = vg.GpuRunner()
pat = vg.Pattern(pat_s)
span = r.batch_size * wps
for it in range(3):
    a = (1 << 64) + (1 << 45) + it * span
    t0 = time.perf_counter(); r.set_pattern(pat, fmt)
    t1 = time.perf_counter(); _, scanned = r.scan_raw(a, a + span - 1, max_hits=0, hits_cap=1, fetch_all=False)
    t2 = time.perf_counter(); total = int(r._lib.vgen_cuda_scan_total_hits(r._ctx)); recs = r.fetch_records(total)
    t3 = time.perf_counter(); m = [vg.GeneratedAddress(a, w, k, fmt) for k, a, w in zip(*recs)]
    t4 = time.perf_counter()
    print("iter %d: set_pattern %.1f ms, scan %.1f ms, fetch_records %.1f ms, objects %.1f ms, hits %d, keys %d, scan-only %.0f Mkeys/s" %
          (it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, total, scanned, scanned / (t2 - t1) / 1e6))
r.close()
