"""The `vgen` CLI (vgen_b200/csrc/cli/vgen.cpp), mirroring the reference's commands (src/lib.rs:44-211). CPU tests cover
argument handling and the no-GPU failure mode; GPU tests compare command output with the oracle."""
import csv
import io
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VGEN = os.path.join(ROOT, "vgen_b200", "vgen")


def run(*args, ok=True):
    p = subprocess.run([VGEN, *args], capture_output=True, text=True, timeout=600)
    if ok:
        assert p.returncode == 0, p.stderr
    return p


@pytest.fixture(scope="module", autouse=True)
def built():
    if not os.path.exists(VGEN):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "vgen_b200", "csrc"), "-s", "-j8"])


def test_usage_and_argument_errors():
    assert "generate" in run("--help").stdout and "list-gpus" in run("--help").stdout
    assert run(ok=False).returncode == 2 and run("frobnicate", ok=False).returncode == 2
    p = run("generate", ok=False)
    assert p.returncode == 1 and "--pattern" in p.stderr
    p = run("range", "-r", "ff:1", ok=False)
    assert p.returncode == 1 and "below" in p.stderr                       # the reference underflows (src/scanner.rs:228)
    assert "START:END" in run("range", "-r", "12", ok=False).stderr
    assert "between 1 and 160" in run("range", "--puzzle", "161", ok=False).stderr
    assert "possible values" in run("range", "-r", "1:2", "-f", "p2xx", ok=False).stderr
    assert "Either --range" in run("range", "-p", "^1A", ok=False).stderr
    assert "Invalid key format" in run("verify", "-k", "nonsense", ok=False).stderr
    assert "not in the embedded table" in run("range", "-p", "boha:b1000:67", ok=False).stderr


def test_no_cpu_fallback():
    # north_star: no CPU fallback. --no-gpu is refused outright; without a GPU every scanning command fails loudly.
    p = run("range", "-r", "1:ff", "--no-gpu", ok=False)
    assert p.returncode == 1 and "no CPU scan path" in p.stderr
    import torch
    if not torch.cuda.is_available():
        p = run("range", "-r", "1:ff", ok=False)
        assert p.returncode == 1 and "GPU initialization failed" in p.stderr
        assert "(none)" in run("list-gpus").stdout and json.loads(run("list-gpus", "--json").stdout) == []


# ---------------------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_range_outputs_match_oracle(tmp_path):
    import oracle
    start, n = 2**64, 1 << 20
    keys, ops, _ = oracle.scan_range("p2pkh", "^1Ab", start, n)
    want = [oracle.address("p2pkh", k) for k in keys]
    rng = "%x:%x" % (start, start + n - 1)
    # jsonl: one record per match, reference field order (src/lib.rs:244-254)
    out = run("range", "-r", rng, "-p", "^1Ab", "-c", "0", "-o", "jsonl", "-q").stdout
    recs = [json.loads(l) for l in out.splitlines()]
    assert [list(r.keys()) for r in recs][:1] == [["address", "wif", "private_key_hex", "format", "pattern", "operations", "elapsed_secs", "rate"]]
    assert [(r["address"], r["wif"], int(r["private_key_hex"], 16)) for r in recs] == [(a, w, k) for (a, w, _), k in zip(want, keys)]
    assert all(r["format"] == "P2PKH" and r["pattern"] == "^1Ab" and r["operations"] == ops for r in recs)
    # csv with header, minimal = WIFs, text blocks, json pretty
    rows = list(csv.reader(io.StringIO(run("range", "-r", rng, "-p", "^1Ab", "-c", "0", "-o", "csv", "-q").stdout)))
    assert rows[0] == ["address", "wif", "private_key_hex", "format", "pattern", "operations", "elapsed_secs", "rate"] and len(rows) == len(keys) + 1
    assert run("range", "-r", rng, "-p", "^1Ab", "-c", "0", "-o", "minimal", "-q").stdout.split() == [w for _, w, _ in want]
    txt = run("range", "-r", rng, "-p", "^1Ab", "-c", "2").stdout
    assert "=== Match 1 of 2 ===" in txt and "Address : " + want[0][0] in txt and "Ops     :" in txt
    f = tmp_path / "o.json"
    p = run("range", "-r", rng, "-p", "^1Ab", "-o", "json", "--file", str(f))           # default count = 1
    assert json.loads(f.read_text())["address"] == want[0][0] and "Wrote 1 result(s)" in p.stderr
    # default pattern `.` and default count 1 -> the first valid key (src/lib.rs:519)
    assert run("range", "-r", "0:ff", "-o", "minimal", "-q").stdout.strip() == oracle.address("p2pkh", 1)[1]
    assert "No match found after 255 operations" in run("range", "-r", "1:ff", "-p", "^1ZZZZZZ", "-c", "0").stderr


@pytest.mark.gpu
def test_provider_puzzle_and_formats():
    import oracle
    out = run("range", "-p", "boha:b1000:20", "-o", "jsonl", "-q")
    rec = json.loads(out.stdout)
    assert int(rec["private_key_hex"], 16) == 0xd2c55 and rec["address"] == "1HsMJxNiV7TLxmoF6uJNkydxPFDog4NQum"
    assert "exact match" in run("range", "-p", "boha:b1000:5", "-q").stderr
    for fmt, pat in (("p2wpkh", "^bc1qq"), ("p2tr", "^bc1pq"), ("ethereum", "^0xab"), ("p2sh-p2wpkh", "^3A")):
        keys, ops, _ = oracle.scan_range(fmt, pat, 2**70, 1 << 18)
        recs = [json.loads(l) for l in run("range", "-r", "%x:%x" % (2**70, 2**70 + (1 << 18) - 1), "-p", pat, "-f", fmt, "-c", "0", "-o", "jsonl", "-q").stdout.splitlines()]
        assert [int(r["private_key_hex"], 16) for r in recs] == keys and [r["address"] for r in recs] == [oracle.address(fmt, k)[0] for k in keys]


@pytest.mark.gpu
def test_verify_estimate_list_generate():
    import oracle
    out = run("verify", "-k", "0c28fca386c7a227600b2fe50b7cae11ec86d3bf1fbe471be89827e19d72aa1d", "-a", "1GAehh7TsJAHuUAeKZcXf5CnwuGuGgyX2S").stdout
    k = 0x0c28fca386c7a227600b2fe50b7cae11ec86d3bf1fbe471be89827e19d72aa1d
    assert "WIF (uncompr.):     5HueCGU8rMjxEXxiPuD5BDku4MkFqeZyd4dZ1jvhTVqvbTLvyTJ" in out and "MATCH!" in out      # README.md:121-127
    for label, fmt in (("P2PKH address:      ", "p2pkh"), ("P2PKH (uncompr.):   ", "p2pkh-uncompressed"), ("P2WPKH address:     ", "p2wpkh"),
                       ("P2SH-P2WPKH addr:  ", "p2sh-p2wpkh"), ("P2TR address:       ", "p2tr"), ("Ethereum address:   ", "ethereum")):
        assert label + oracle.address(fmt, k)[0] in out
    assert "MISMATCH!" in run("verify", "-k", "5HueCGU8rMjxEXxiPuD5BDku4MkFqeZyd4dZ1jvhTVqvbTLvyTJ", "-a", "1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH").stdout
    eth = oracle.address("ethereum", k)[0]
    assert "case-insensitive" in run("verify", "-k", "%064x" % k, "-a", eth.lower()[2:]).stdout
    est = run("estimate", "-p", "^1Cat").stdout
    assert "Estimated difficulty: 1 in 195112" in est and "Benchmark rate:" in est
    assert "NVIDIA" in run("list-gpus").stdout and json.loads(run("list-gpus", "--json").stdout)[0]["backend"] == "Cuda"
    rec = json.loads(run("generate", "-p", "^1Ab", "-o", "jsonl", "-q").stdout)
    assert rec["address"].startswith("1Ab") and oracle.address("p2pkh", int(rec["private_key_hex"], 16))[0] == rec["address"]
    p = run("range", "-r", "1:ff", "-p", "^1O", "-q")                        # impossible pattern: lint warning (src/lib.rs:686-708)
    assert "NEVER match" in p.stderr and "Base58 excludes" in p.stderr


@pytest.mark.gpu
def test_checkpoint_resume(tmp_path):
    import oracle
    start, n = 2**66, 5 << 20
    keys, _, _ = oracle.scan_range("p2pkh", "^1A", start, n)
    ck, out = tmp_path / "ck", tmp_path / "out.jsonl"
    args = ["range", "-r", "%x:%x" % (start, start + n - 1), "-p", "^1A", "-c", "0", "-o", "jsonl", "-q", "--file", str(out), "--checkpoint", str(ck), "--checkpoint-keys", str(1 << 20)]
    run(*args)
    got = [int(json.loads(l)["private_key_hex"], 16) for l in out.read_text().splitlines()]
    assert got == keys
    assert "next=%064x" % (start + n) in ck.read_text()
    # roll the mark back to the 3rd segment and drop the later output: a resumed run completes the set without redoing the first two
    txt = ck.read_text().splitlines()
    cut = start + (2 << 20)
    kept = "".join(l + "\n" for l in out.read_text().splitlines() if int(json.loads(l)["private_key_hex"], 16) < cut)
    ck.write_text("\n".join(l if not l.startswith(("next=", "found=", "operations=", "out_bytes=")) else
                            {"ne": "next=%064x" % cut, "fo": "found=%d" % sum(k < cut for k in keys), "op": "operations=%d" % (2 << 20),
                             "ou": "out_bytes=%d" % len(kept)}[l[:2]] for l in txt) + "\n")
    out.write_text(kept)
    run(*args)
    assert [int(json.loads(l)["private_key_hex"], 16) for l in out.read_text().splitlines()] == keys
    # a crash between "matches written" and "checkpoint renamed": the output holds matches beyond the mark. The resumed run cuts the
    # file back to the checkpointed size before it scans those keys again, so nothing is reported twice.
    full = out.read_text()
    ck.write_text("\n".join(l if not l.startswith(("next=", "found=", "operations=", "out_bytes=")) else
                            {"ne": "next=%064x" % cut, "fo": "found=%d" % sum(k < cut for k in keys), "op": "operations=%d" % (2 << 20),
                             "ou": "out_bytes=%d" % len(kept)}[l[:2]] for l in txt) + "\n")
    assert len(full) > len(kept)
    run(*args)
    assert [int(json.loads(l)["private_key_hex"], 16) for l in out.read_text().splitlines()] == keys
    # a hit limit reached inside a segment: the mark moves to the last match + 1, so a later run with a larger limit continues there
    ck2, out2 = tmp_path / "ck2", tmp_path / "out2.jsonl"
    lim = [a for a in args]
    lim[lim.index("-c") + 1] = "5"; lim[lim.index(str(out))] = str(out2); lim[lim.index(str(ck))] = str(ck2)
    run(*lim)
    assert [int(json.loads(l)["private_key_hex"], 16) for l in out2.read_text().splitlines()] == keys[:5]
    assert "next=%064x" % (keys[4] + 1) in ck2.read_text() and "found=5" in ck2.read_text()
    lim[lim.index("-c") + 1] = "8"
    run(*lim)
    assert [int(json.loads(l)["private_key_hex"], 16) for l in out2.read_text().splitlines()] == keys[:8]
    # a checkpoint of another scan is refused
    p = run(*[a if a != "^1A" else "^1B" for a in args], ok=False)
    assert p.returncode == 1 and "different scan" in p.stderr
