"""tests/cabi_harness.c is a plain-C client of include/vgen_cuda.h that makes the call sequence of the Rust binding
(src/gpu.rs): create -> set_pattern -> scan with a progress callback and a stop flag -> fetch_hits -> format -> frames ->
destroy. No Rust toolchain exists in this image, so this is the compiled, non-Python check of the drop-in boundary.
CPU part: the header is valid C99 and the harness links against the library. GPU part: its output equals the oracle's."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cabi_harness.c")
EXE = os.path.join(ROOT, "tests", "host_emu", "cabi_harness")
LIBDIR = os.path.join(ROOT, "vgen_b200")


def build():
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(SRC), os.path.getmtime(os.path.join(ROOT, "include", "vgen_cuda.h"))):
        subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-O1", "-I" + os.path.join(ROOT, "include"), SRC, "-o", EXE,
                               "-L" + LIBDIR, "-lvgen_cuda", "-Wl,-rpath," + LIBDIR])
    return EXE


def test_harness_compiles_as_c99_and_links():
    exe = build()
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 1 and "usage" in out.stderr          # no GPU needed for the usage path


def test_rust_binding_declares_what_the_header_declares():
    """src/gpu.rs (uncompilable here) must at least agree with the header on every function it binds: same name, same arity."""
    hdr = open(os.path.join(ROOT, "include", "vgen_cuda.h")).read()
    rs = open(os.path.join(ROOT, "src", "gpu.rs")).read()
    block = rs[rs.index('extern "C" {'):]
    block = block[:block.index("\n}\n")]
    decls = re.findall(r"fn (vgen_\w+)\s*\(([^;]*?)\)\s*(?:->\s*[^;]+)?;", block, re.S)
    assert len(decls) >= 14
    for name, args in decls:
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, hdr, re.S)
        assert m, name + " is not in include/vgen_cuda.h"
        c_args = [a for a in m.group(1).split(",") if a.strip() and a.strip() != "void"]
        rs_args = [a for a in args.split(",") if a.strip()]
        assert len(c_args) == len(rs_args), (name, c_args, rs_args)
    # every pub item of the reference's gpu.rs that callers use is still there
    for item in ("pub enum GpuBackend", "pub fn fallback_order", "pub fn name", "pub const DEFAULT_BATCH_SIZE", "pub struct GpuRunner", "pub batch_size: u32",
                 "pub device_name: String", "pub fn backend(&self)", "pub async fn new(batch_size: u32, backend: GpuBackend)", "pub fn dispatch(&self, start_key: [u8; 32], frame_index: usize)",
                 "pub async fn await_result(&self, frame_index: usize) -> Result<(Vec<[u8; 20]>, [u8; 32])>", "pub async fn generate_batch(&self, start_key: [u8; 32]) -> Result<Vec<[u8; 20]>>",
                 "pub fn dispatch_p2tr(&self, start_key: [u8; 32], frame_index: usize)", "pub async fn await_result_p2tr(&self, frame_index: usize) -> Result<(Vec<[u8; 32]>, [u8; 32])>",
                 "pub fn dispatch_batch_hash(&self, start_key: [u8; 32], frame_index: usize)", "pub async fn scan_gpu_with_runner(", "pub fn scan_gpu(",
                 "pub async fn scan_gpu_p2tr_with_runner(", "pub fn scan_gpu_p2tr(", "pub(crate) fn is_software_adapter("):
        assert item in rs, item


def _run(*args):
    out = subprocess.run([build()] + [str(a) for a in args], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr
    rec = {"HIT": []}
    for line in out.stdout.splitlines():
        k, _, rest = line.partition(" ")
        if k == "HIT":
            rec["HIT"].append(rest.split())
        else:
            rec[k] = rest
    assert "DONE" in rec
    return rec


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,fid,pat", [("p2pkh", 0, "^1Ca"), ("p2wpkh", 2, "qq$"), ("ethereum", 5, "^0xAb"), ("p2tr", 4, "^bc1pqq")])
def test_harness_scan_equals_oracle(fmt, fid, pat):
    import oracle
    start, n = 2**64, 1 << 20 if fmt != "p2tr" else 1 << 18
    rec = _run(fid, pat, "%x" % start, n, 0)
    keys, ops, _ = oracle.scan_range(fmt, pat, start, n)
    assert len(keys) > 4                                            # more hits than the harness's caller buffer holds
    assert [int(h[0], 16) for h in rec["HIT"]] == keys
    for h in rec["HIT"]:
        a, w, _ = oracle.address(fmt, int(h[0], 16))
        assert (h[1], h[2]) == (a, w)
    toks = ("SCANNED " + rec["SCANNED"]).split()
    f = dict(zip(toks[0::2], toks[1::2]))
    assert int(f["SCANNED"]) == ops == int(f["LAST"]) and int(f["TOTAL"]) == len(keys) and int(f["FIRST"]) == 4
    assert int(f["CALLBACKS"]) >= 1 and f["MONOTONE"] == "1" and f["STOPPED"] == "0"          # one callback per window
    assert "regex" in rec["BADPATTERN"].lower() or "pattern" in rec["BADPATTERN"].lower()
    # the frame calls: hash160 of consecutive keys, frame 1 continuing where frame 0 ended; P2TR frame = untweaked x-only keys
    assert bytes.fromhex(rec["FRAME0"]) == oracle.hash160_batch(start, 4096)
    assert bytes.fromhex(rec["FRAME1"]) == oracle.hash160_batch(start + 4096, 4096)
    xs = bytes.fromhex(rec["FRAMEX"])
    for i in range(8):
        assert int.from_bytes(xs[32 * i:32 * i + 32], "big") == oracle.pubkey(start + i)[0]


@pytest.mark.gpu
def test_harness_stop_flag_and_hit_limit():
    import oracle
    start = 2**64
    # stop flag raised from inside the second progress callback: the scan ends early, counts stay consistent
    rec = _run(0, "^1CatDogs", "%x" % start, 1 << 30, 0, 2)
    toks = ("SCANNED " + rec["SCANNED"]).split()
    f = dict(zip(toks[0::2], toks[1::2]))
    assert f["STOPPED"] == "1" and 0 < int(f["SCANNED"]) < (1 << 30) and int(f["SCANNED"]) == int(f["LAST"])
    # hit limit: the first 3 matches in ascending key order
    rec = _run(0, "^1A", "%x" % start, 1 << 20, 3)
    keys, _, _ = oracle.scan_range("p2pkh", "^1A", start, 1 << 16)
    assert [int(h[0], 16) for h in rec["HIT"]] == keys[:3]
