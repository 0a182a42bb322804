#!/usr/bin/env python
"""bench.py — throughput of the B200 key-scan hot path (Mkeys/s, P2PKH ^1Cat prefix scan) with its integer-issue
roofline, the end-to-end figure through the public API, and the CPU baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W]            own arm (one process per GPU under torchrun for N > 1)
  python bench.py --single-context --gpus N [...]                own arm, ONE process and one context over N GPUs (what
                                                                 `vgen --devices 0,..,N-1` and the Rust binding use; hits merged on the host)
  python bench.py --single-context --gpus N --c5                 BASELINE config C5: the fixed 2^40-key exact-address window (strong scaling)
  python bench.py --impl reference [--gpus N] [--steps K] ...    the reference's CPU scan path (oracle port) on host cores

A step = one pass of the scan path over `--windows-per-step` full launch windows per GPU (fixed synthetic key range
from 2^64; keys are a counter, nothing is loaded). Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Mkeys/s P2PKH ^1Cat scan"
START = 1 << 64
SM_COUNT, LANES_PER_SM = 148, 64
# algorithmic work per key, 32-bit integer instruction slots (DESIGN.md "Roofline"; SURVEY.md 8d, +/- variant, c_w = 1):
#   EC 3.5M + 1S = 300.5 limb products + ~300 carry/add slots, SHA-256 ~1400, RIPEMD-160 ~975, matcher ~15
W_INT = {"p2pkh": 3050, "p2wpkh": 3050, "p2sh-p2wpkh": 5425, "p2pkh-uncompressed": 4450,
         # Ethereum: Keccak-f[1600] on 32-bit halves = 24 x (122 LOP3 + 58 SHF) = 4,320 + walk (both coordinates) ~1,080;
         # ncu counts 5,509 executed per key (profiles/ncu_walk_eth_r01_summary.txt)
         "ethereum": 5400,
         # P2TR (round 2: output key from the key's scalar, XYZZ accumulator, 9-window signed table that bench_resident builds as the
         # long-job setting): walk 4.5 field mults + tweak stage 4M+2S, 6 x (8M+2S), 5M+2S over 9 table entries + 4 + 270/64 mults for
         # the grouped inversion = ~90 field mults x ~147 slots + one SHA-256 compression (round 1, 10 Jacobian additions: 19,400)
         "p2tr": 14700}
HBM_BYTES_PER_KEY = 32          # prefix-product scratch: 32 B written + 32 B read per +/- key pair
# BASELINE config C5: P2PKH address of the key at offset int(0.83 * 2^40) of the window from 2^65 (derived with the oracle by
# tests/test_full_size_configs.py::test_c5_exact_address_2_40, which fails if this constant is wrong)
C5_START, C5_KEYS, C5_OFFSET, C5_ADDRESS = 1 << 65, 1 << 40, 912594651054, "1D6vERHvZo8isjfDVdYXkoCtuMYxgaZEVG"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"], "samples": 0}
        sm = sorted(float(r[1]) for r in rows)
        reasons = []
        for i, name in ((4, "hw_slowdown"), (5, "hw_thermal_slowdown"), (6, "sw_thermal_slowdown"), (7, "sw_power_cap")):
            if any(r[i].lower().startswith("active") for r in rows):
                reasons.append(name)
        pw = [float(r[3]) for r in rows if r[3].replace(".", "", 1).isdigit()]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(pw) if pw else None}


def load_json(path):
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def measure_int_peaks(device_index):
    """Runs vgen_b200/peak_bench (integer issue-rate microbenchmarks, < 1 s) on this GPU inside the bench run, with the SM clock
    sampled while it runs. Returns (peaks dict, clock summary) or (None, reason)."""
    exe = os.path.join(ROOT, "vgen_b200", "peak_bench")
    if not os.path.exists(exe):
        return None, "vgen_b200/peak_bench is not built"
    sampler = ClockSampler(device_index)
    sampler.start()
    t0 = time.perf_counter()
    try:
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(device_index))
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120, env=env)
        peaks = json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:       # the bench line then says so instead of quoting an old file
        sampler.stop()
        return None, "peak_bench failed: %r" % (e,)
    t1 = time.perf_counter()
    sampler.stop()
    return peaks, sampler.summary(t0, t1)


def cpu_scan_sample(fmt, pattern, threads, target_s):
    """Times the oracle's scan_range_cpu port on a bounded sample of the same workload (same start, format, pattern)."""
    import oracle
    n = 1 << 12
    for _ in range(2):          # calibrate twice (thread start-up dominates the first tiny run)
        n *= 16
        keys, ops, el = oracle.scan_range(fmt, pattern, START, n, threads=threads)
        rate = ops / max(el, 1e-6)
    n = max(1 << 14, min(1 << 26, int(rate * target_s)))
    keys, ops, el = oracle.scan_range(fmt, pattern, START, n, threads=threads)
    return ops, el, len(keys)


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (scan_range_cpu + AddressGenerator::generate,
    src/scanner.rs:211-330) as restated by oracle/ (the Rust reference cannot be built here: no cargo/rustc), on all host threads."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    import oracle
    oracle.lib()
    per_step_s = 3.0
    for _ in range(max(args.warmup, 1)):
        cpu_scan_sample(args.format, args.pattern, threads, 0.5)
    tot_ops, tot_el, sample_keys = 0, 0.0, 0
    for _ in range(args.steps):
        ops, el, _ = cpu_scan_sample(args.format, args.pattern, threads, per_step_s)
        tot_ops += ops
        tot_el += el
        sample_keys = ops
    val = tot_ops / tot_el / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Mkeys/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": tot_el / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
        "data": "synthetic", "config": workload_config(args),
        "cpu_baseline": {"value": val, "unit": "Mkeys/s", "cores": threads, "kind": "port",
                         "sample": "%d consecutive keys from 2^64 per step (~%.0f s), %s %s, %d steps" % (sample_keys, per_step_s, args.format, args.pattern, args.steps)},
        "e2e": {"value": val, "unit": "Mkeys/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "oracle port of vgen's rayon CPU scanner (one independent k*G per key, Base58Check + regex per key); README.md:175 quotes 0.05-0.2 Mkeys/s",
    }
    print(json.dumps(line))
    return 0


def workload_config(args):
    """The workload both arms run (identical dict in the own arm and the reference arm; what the GPU arm derives from the
    device — window size, matcher kind — is reported under the separate `geometry` key)."""
    if args.format == "p2pkh" and args.pattern == "^1Cat":
        what = "BASELINE configs[1] shape: P2PKH compressed prefix scan (hash160 interval-compare path), pattern ^1Cat, fixed key range from 2^64"
    else:
        what = "%s scan, pattern %s, fixed key range from 2^64 (not the headline workload)" % (args.format, args.pattern)
    return {"workload": what,
            "format": args.format, "pattern": args.pattern, "start": "2^64", "windows_per_step": args.windows_per_step,
            "parallelism": "keyspace shards over %d GPU(s), no collective; 2 overlapping streams per GPU" % args.gpus,
            "l2": "per-launch working set (prefix-product scratch, 1.24 GB) exceeds the 126 MB L2; no flush needed"}


def roofline_block(args, runner, ms, n_launches, keys, clocks, n_gpus_timed):
    """Roofline of the walk kernel (the only kernel in the timed region): INT32 instruction issue, not HBM / tensor.
    achieved = algorithmic instruction slots per key (W_INT) x keys per launch / the launch duration measured in THIS run
    (CUDA events over the timed region / launches per GPU). peak = nominal 148 SMs x 64 lanes x max SM clock; the pipe peaks
    measured by peak_bench in this same run are reported beside it."""
    peaks = load_json(os.path.join(ROOT, "MEASURED_PEAKS.json")) or {}
    ncu = load_json(os.path.join(ROOT, "profiles", "walk_kernel_ncu.json")) or {}
    int_peaks, peak_clocks = measure_int_peaks(env_int("LOCAL_RANK", 0))
    sm_max = clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    sm_now = clocks.get("sm_mhz") or sm_max
    w_int = W_INT.get(args.format, 3050)
    keys_per_gpu_s = keys / n_gpus_timed / (ms * 1e-3)
    achieved = keys_per_gpu_s * w_int / 1e12
    nominal_peak = SM_COUNT * LANES_PER_SM * sm_max * 1e6 / 1e12
    launches_per_gpu = n_launches / n_gpus_timed
    out = {"bound": "int32-issue", "achieved": achieved, "peak": nominal_peak, "unit": "Tinst/s", "frac": achieved / nominal_peak,
           "peak_source": "nominal: %d SMs x %d INT32 lanes x %.0f MHz max SM clock (the '148-SM IMAD issue peak' of BASELINE.json)" % (SM_COUNT, LANES_PER_SM, sm_max),
           "frac_note": "literal fraction of the 64-lane IMAD issue peak; multiplies (FMA-heavy pipe) and add/logic/shift (ALU pipe) issue on two 64-lane pipes, so it can exceed 1",
           "frac_at_sampled_clock": achieved / (SM_COUNT * LANES_PER_SM * sm_now * 1e6 / 1e12),
           "work_per_key_inst": w_int, "keys_per_launch": runner.batch_size,
           "launch_ms": ms / launches_per_gpu,
           "launch_ms_note": "CUDA-event time of the timed region / walk-kernel launches per GPU = effective duration of one launch. A launch fills every "
                             "CTA slot of the GPU (296 CTAs); the %d lanes (streams) only overlap the tail of one window with the head of the next, so one "
                             "kernel in isolation lasts ~5%% longer than this (ncu, serialised: profiles/)" % (runner.lane_count() // max(1, len(runner.devices))),
           "executed_inst_per_key_ncu": ncu.get("thread_inst_per_key"),
           "traffic": ncu.get("dram_bytes_per_launch"),
           "ncu_source": {k: ncu.get(k) for k in ("source", "commit", "date", "kernel", "command")} if ncu else None,
           "hbm": {"achieved": keys_per_gpu_s * HBM_BYTES_PER_KEY / 1e9, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                   "frac": (keys_per_gpu_s * HBM_BYTES_PER_KEY / 1e9) / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
                   "peak_source": "of measured (MEASURED_PEAKS.json)" if peaks.get("hbm_gbs") else "absent"}}
    if int_peaks:
        # the binding unit is the ALU pipe: SHA-256 + RIPEMD-160 are add/logic/shift work. ALU-pipe slots per key: measured by
        # ncu (sm__inst_executed_pipe_alu) when a capture exists, else the model (W_INT minus the wide products)
        alu_slots = ncu.get("alu_pipe_inst_per_key") or (w_int - 360)
        alu_peak = min(int_peaks["alu_lop3_shf"], int_peaks["alu_iadd3"])
        out["measured_pipe_peaks"] = dict(int_peaks, clocks_during_peak_bench=peak_clocks, how="vgen_b200/peak_bench run inside this bench run; SASS instruction counts checked by tests/test_peak_bench_sass.py")
        out["binding_unit"] = {"unit": "alu-pipe", "slots_per_key": alu_slots, "achieved": keys_per_gpu_s * alu_slots / 1e12, "peak": alu_peak,
                               "frac": keys_per_gpu_s * alu_slots / 1e12 / alu_peak, "peak_source": "measured in this run: min(alu_lop3_shf, alu_iadd3)"}
        out["frac_of_measured_imad_peak"] = achieved / int_peaks["imad"]
        out["issue_mix_1_3"] = {"achieved": achieved, "peak": int_peaks["mix_imad_alu_1_3"], "frac": achieved / int_peaks["mix_imad_alu_1_3"]}
    else:
        out["measured_pipe_peaks"] = None
        out["measured_pipe_peaks_note"] = str(peak_clocks)
    return out


def run_c5(args):
    """BASELINE config C5 as written: exact-address match over the fixed 2^40-key window from 2^65, one context over N GPUs
    (strong scaling: the window is fixed, the time shrinks). The planted key is the only hit."""
    import vgen_b200 as vg
    n = args.gpus
    runner = vg.GpuRunner(devices=list(range(n)))
    pat = vg.Pattern("^" + C5_ADDRESS + "$")
    sampler = ClockSampler(0)
    sampler.start()
    warm = vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=vg.UNLIMITED, start=C5_START - (1 << 34), end=C5_START - 1)
    vg.scan_gpu_with_runner(pat, warm, None, None, runner)
    io0 = runner.io_bytes()
    l0 = runner.launch_count()
    t0 = time.perf_counter()
    res = vg.scan_gpu_with_runner(pat, vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=vg.UNLIMITED, start=C5_START, end=C5_START + C5_KEYS - 1), None, None, runner)
    t1 = time.perf_counter()
    sampler.stop()
    io1 = runner.io_bytes()
    ok = [int(m.hex, 16) for m in res.matches] == [C5_START + C5_OFFSET] and res.matches[0].address == C5_ADDRESS and res.operations == C5_KEYS
    secs = t1 - t0
    line = {"metric": "seconds for the 2^40-key exact-address window (BASELINE config C5)", "value": secs, "unit": "s", "n_gpus": n, "steps": 1, "warmup": 1,
            "ms_per_step": secs * 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[4]: P2PKH exact-address match over the fixed window [2^65, 2^65 + 2^40), planted key at offset int(0.83 * 2^40)",
                       "pattern": "^" + C5_ADDRESS + "$", "parallelism": "one context, %d GPU(s), one host thread per GPU, hits merged on the host; no collective" % n,
                       "l2": "per-launch working set (1.24 GB) exceeds the 126 MB L2"},
            "mkeys_per_s": C5_KEYS / secs / 1e6, "planted_hit_found_and_alone": ok, "clocks": sampler.summary(t0, t1),
            "e2e": {"value": secs, "unit": "s", "h2d_bytes_per_step": io1[0] - io0[0], "d2h_bytes_per_step": io1[1] - io0[1],
                    "api": "vgen_b200.scan_gpu_with_runner (one vgen_cuda_scan call over the whole window)"},
            "gpu_launches": runner.launch_count() - l0, "matcher": runner.matcher_kind(), "stats": runner.stats(), "device": runner.device_name}
    runner.close()
    print(json.dumps(line))
    return 0 if ok else 1


def run_own(args):
    import vgen_b200 as vg
    single = args.single_context
    if single:
        rank, world, local = 0, 1, 0
        n_ctx_gpus = args.gpus
    else:
        import torch
        import torch.distributed as dist
        rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
        n_ctx_gpus = 1
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the scan path has no CPU fallback")
        torch.cuda.set_device(local)
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dev = torch.device("cuda", local)

    def barrier():
        if single:
            return
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if single:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if single:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    if single:
        runner = vg.GpuRunner(devices=list(range(n_ctx_gpus)))
    else:
        runner = vg.GpuRunner(devices=[local], shard=(rank, world))
    fmt = vg.AddressFormat.parse(args.format)
    pat = vg.Pattern(args.pattern)
    runner.set_pattern(pat, fmt)
    wps, K, W = args.windows_per_step, args.steps, args.warmup
    step_keys = runner.batch_size * wps                 # per GPU

    # ---- device-resident arm: centres already in HBM when the timed region starts; CUDA events on the launch streams
    # (per GPU; the slowest GPU's time counts)
    runner.bench_resident(START, max(W, 3) * wps)       # warm-up
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    barrier()
    l0 = runner.launch_count()
    t0 = time.perf_counter()
    ms, keys, launches = runner.bench_resident(START + (1 << 40), K * wps)
    barrier()
    t1 = time.perf_counter()
    gpu_launches = runner.launch_count() - l0 - runner.lane_count()        # minus the untimed centre set-up launches (one per lane)
    ms_max = allmax(ms)
    total_keys = allsum(keys)
    value = total_keys / (ms_max * 1e-3) / 1e6
    clocks = sampler.summary(t0, t1)

    # ---- end-to-end arm: the public call (scan_gpu_with_runner) with host-side start/end scalars, centre set-up,
    # launch parameters, hit records and counters crossing PCIe inside the timed region
    n_total_gpus = world * n_ctx_gpus
    span = step_keys * n_total_gpus
    cfg0 = vg.ScanConfig(format=fmt, count=vg.UNLIMITED, start=START + (1 << 44), end=START + (1 << 44) + span - 1)
    vg.scan_gpu_with_runner(pat, cfg0, None, None, runner)          # warm-up
    io0 = runner.io_bytes()
    barrier()
    e0 = time.perf_counter()
    e_keys, e_hits = 0, 0
    for s in range(K):
        a = START + (1 << 45) + s * span
        res = vg.scan_gpu_with_runner(pat, vg.ScanConfig(format=fmt, count=vg.UNLIMITED, start=a, end=a + span - 1), None, None, runner)
        e_keys += res.operations
        e_hits += len(res.matches)
    barrier()
    e_ms = allmax((time.perf_counter() - e0) * 1e3)
    sampler.stop()
    io1 = runner.io_bytes()
    e_total = allsum(e_keys)
    e_hits_total = allsum(e_hits)
    e2e = {"value": e_total / (e_ms * 1e-3) / 1e6, "unit": "Mkeys/s",
           "h2d_bytes_per_step": (io1[0] - io0[0]) // K, "d2h_bytes_per_step": (io1[1] - io0[1]) // K,
           "ms_per_step": e_ms / K, "hits_per_step": e_hits_total / K,
           "api": "vgen_b200.scan_gpu_with_runner -> vgen_cuda_scan + vgen_cuda_scan_fetch_text (host scalars in; confirmed hits formatted as key/address/WIF text and copied to the caller; Python objects are built lazily on access)"}

    line = {"metric": METRIC, "value": value, "unit": "Mkeys/s", "n_gpus": n_total_gpus, "steps": K, "warmup": max(W, 3), "ms_per_step": ms_max / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": workload_config(args), "clocks": clocks, "e2e": e2e, "gpu_launches": int(gpu_launches),
            "geometry": {"keys_per_window": runner.batch_size, "keys_per_step_per_gpu": step_keys, "matcher": runner.matcher_kind(),
                         "lanes": runner.lane_count(), "context": "one context over %d GPU(s), one host thread per GPU" % n_ctx_gpus if single else "one process per GPU (torchrun), %d rank(s)" % world,
                         "stats": runner.stats()},
            "device": runner.device_name}
    if rank == 0:
        line["roofline"] = roofline_block(args, runner, ms, launches, keys, clocks, n_ctx_gpus)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        ops, el, nh = cpu_scan_sample(args.format, args.pattern, threads, 12.0)
        line["cpu_baseline"] = {"value": ops / el / 1e6, "unit": "Mkeys/s", "cores": threads, "kind": "port",
                                "sample": "first %d keys of the same range (from 2^64), %s %s, %.1f s, %d matches" % (ops, args.format, args.pattern, el, nh)}
    runner.close()
    if not single and world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--format", default="p2pkh")
    ap.add_argument("--pattern", default="^1Cat")
    ap.add_argument("--windows-per-step", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--single-context", action="store_true", help="one process, one context over --gpus devices (no torchrun)")
    ap.add_argument("--c5", action="store_true", help="with --single-context: time BASELINE config C5 (fixed 2^40-key exact-address window)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.c5:
        return run_c5(args)
    world = env_int("WORLD_SIZE", 1)
    if args.gpus > 1 and world == 1 and not args.single_context:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_own(args)


if __name__ == "__main__":
    sys.exit(main())
