"""Experiment: several lanes (streams with their own centres / scratch) on ONE GPU, so that the CTAs of the next window
fill the slots that the previous window's early finishers leave empty. Prints Mkeys/s through the public scan call."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgen_b200 as vg

pat = vg.Pattern(sys.argv[1] if len(sys.argv) > 1 else "^1CatDog")
fmt = vg.AddressFormat.parse(sys.argv[2] if len(sys.argv) > 2 else "p2pkh")
nwin = int(sys.argv[3]) if len(sys.argv) > 3 else 48
for devs in ([0], [0, 0], [0, 0, 0]):
    r = vg.GpuRunner(devices=devs)
    span = r.batch_size * nwin
    best = 0
    for it in range(3):
        a = (1 << 64) + (1 << 50) + it * span
        t0 = time.perf_counter()
        res = vg.scan_gpu_with_runner(pat, vg.ScanConfig(format=fmt, count=vg.UNLIMITED, start=a, end=a + span - 1), None, None, r)
        dt = time.perf_counter() - t0
        best = max(best, res.operations / dt / 1e6)
    print("lanes on GPU0: %d  window %d keys  best %.0f Mkeys/s  (hits %d)" % (len(devs), r.batch_size, best, len(res.matches)), flush=True)
    r.close()
