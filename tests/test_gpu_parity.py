"""GPU parity tests (run with -m gpu on a B200): the CUDA scan path, called through the C ABI, against the CPU oracle.
Bit-exact: every digest byte of generate_batch, and identical (key, address) match sets for the scan."""
import ctypes as C
import hashlib
import random

import pytest

import oracle
import vgen_b200 as vg

pytestmark = pytest.mark.gpu
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
FORMATS = ["p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "ethereum"]


@pytest.fixture(scope="module")
def runner():
    r = vg.GpuRunner()
    yield r
    r.close()


@pytest.fixture(scope="module")
def small_runner():
    # reference-sized batches (DEFAULT_BATCH_SIZE = 524,288 keys per launch, src/gpu.rs:83)
    r = vg.GpuRunner(vg.DEFAULT_BATCH_SIZE)
    yield r
    r.close()


def test_library_is_the_cuda_one(runner):
    assert "B200" in runner.device_name or runner.device_name
    assert runner.backend() == vg.GpuBackend.Cuda
    assert runner.batch_size % 1024 == 0


def test_generate_batch_golden(runner, golden):
    for r in golden["runs"]:
        got = runner.generate_batch(int(r["start"], 16), len(r["hash160"]))
        want = [h if h else "00" * 20 for h in r["hash160"]]
        assert [g.hex() for g in got] == want, r["start"]


@pytest.mark.parametrize("start", [1, 2**32 - 5, 2**64, N - 2**16 - 3, 0x2832ED74F2B5E35EE - 1000])
def test_hash160_batch_vs_oracle(runner, start):
    n = 1 << 16
    got = b"".join(runner.generate_batch(start, n))
    want = oracle.hash160_batch(start, n)
    assert got == want


def test_hash160_batch_edges(runner):
    # scalar 0 and scalars >= n yield zeros (invalid keys, src/address.rs:93)
    got = runner.generate_batch(0, 5000)
    assert got[0] == b"\0" * 20
    assert b"".join(got[1:]) == oracle.hash160_batch(1, 4999)
    got = runner.generate_batch(N - 3000, 5000)
    assert b"".join(got[:3000]) == oracle.hash160_batch(N - 3000, 3000)
    assert all(g == b"\0" * 20 for g in got[3000:])
    # ragged sizes
    for n in (1, 2, 1023, 1025, 4097):
        assert b"".join(runner.generate_batch(12345, n)) == oracle.hash160_batch(12345, n)
    assert runner.generate_batch(1, 0) == []


@pytest.mark.parametrize("fmt", FORMATS)
def test_digest_batch_all_formats(small_runner, fmt):
    rng = random.Random(42)
    for start in (1, rng.randrange(1, N - 10**6)):
        n = 20000
        got = small_runner.digest_batch(fmt, start, n)
        raw = oracle.digest_batch(fmt, start, n)
        want = b"".join(raw[32 * i:32 * i + 20] for i in range(n))
        assert got == want, (fmt, start)


def test_p2tr_digest_batch_is_the_witness_program(small_runner):
    """P2TR: both stages on the GPU (even-Y internal key, TapTweak, + t*G) == the oracle's output key (src/address.rs:136-140)."""
    rng = random.Random(7)
    for start, n in ((1, 5000), (rng.randrange(1, N - 10**6), 20000), (N - 3000, 4000), (2**64, small_runner.batch_size + 777)):
        got = small_runner.digest_batch("p2tr", start, n)
        want = oracle.digest_batch("p2tr", start, n)
        assert got == want, start


def test_multi_window_walk_consistency(small_runner):
    # 3+ windows of the small runner: the next-centre update (C + W*G) is exercised
    n = small_runner.batch_size * 3 + 777
    start = 2**64 + 99
    got = small_runner.digest_batch("p2pkh", start, n)
    for off in (0, small_runner.batch_size - 5, 2 * small_runner.batch_size - 3, n - 2000):
        m = 2000
        assert got[20 * off:20 * (off + m)] == oracle.hash160_batch(start + off, m)
    assert hashlib.sha256(got).hexdigest() == hashlib.sha256(b"".join(small_runner.generate_batch(start, n))).hexdigest()


def _scan(runner, fmt, pat, start, n, ci=False, count=vg.UNLIMITED):
    cfg = vg.ScanConfig(format=vg.AddressFormat.parse(fmt), count=count, start=start, end=start + n - 1)
    return vg.scan_gpu_with_runner(vg.Pattern(pat, ci), cfg, None, None, runner)


def _check_scan(runner, fmt, pat, start, n, ci=False):
    res = _scan(runner, fmt, pat, start, n, ci)
    keys, ops, _ = oracle.scan_range(fmt, pat, start, n, ci=ci)
    assert [int(m.hex, 16) for m in res.matches] == keys, (fmt, pat)
    assert res.operations == ops
    for m in res.matches:
        a, w, _ = oracle.address(fmt, int(m.hex, 16))
        assert (m.address, m.wif) == (a, w)
    return res


# (format, pattern, case-insensitive, keys): every case has a NON-EMPTY oracle set on [2^64, 2^64 + keys) (asserted), so no
# case compares [] with []; patterns that can never match are in test_scan_patterns_that_cannot_match below.
SMALL_CASES = [
    ("p2pkh", "^1Ca", False), ("p2pkh", "^1cat", True), ("p2pkh", "^1[Cc]a", False), ("p2pkh", "^11", False),
    ("p2pkh-uncompressed", "^1Ab", False), ("p2sh-p2wpkh", "^3Ab", False), ("p2sh-p2wpkh", "^3", False),
    ("p2wpkh", "^bc1qaa", False), ("p2wpkh", "ead$", False), ("p2wpkh", "aa.*zz", False),
    ("ethereum", "^0xdea", False), ("ethereum", "^0xDe", False), ("ethereum", "^0xab", True), ("ethereum", "F00$", False),
    ("p2tr", "^bc1pqq", False), ("p2tr", "ead$", False, 1 << 19), ("p2tr", "^bc1p.*zzz", False),
    ("p2pkh", "abc$", False), ("p2pkh", "[0-9]{4}$", False), ("p2pkh", "^1.*XYZ", False),
    # anchored alternations / optional tails: prefixes read off the DFA become digest intervals (lowering.cpp)
    ("p2pkh", "^(1Ca|1Do)", False), ("p2pkh", "^1(Ca|Do)t?x", False), ("p2wpkh", "^bc1q(aa|zz)", False),
    ("ethereum", "^0x(de|be)a", False), ("p2tr", "^bc1p(qq|pp)", False),
    # Ethereum case-sensitive patterns: the case-folded DFA runs first, the EIP-55 hash only for keys it accepts
    ("ethereum", "[A-F]{3}$", False), ("ethereum", "(?-i:aB)", False), ("ethereum", "e$", False), ("ethereum", "Ab|cD", False),
    # literal suffixes: tested on the Bech32 checksum bits / the last digest bits / the Base58 residue before any DFA step
    ("p2wpkh", "qq$", False), ("p2wpkh", "ZQ$", True), ("p2tr", "7$", False), ("ethereum", "1f$", False), ("ethereum", "AB$", True),
    ("ethereum", "AB$", False), ("ethereum", "07$", False),
    ("p2pkh", "at$", False), ("p2pkh", "AT$", True), ("p2sh-p2wpkh", "Zz$", False), ("p2pkh-uncompressed", "7$", False),
    # the wider regex surface (round 2): word boundaries, flag x, nested classes / set operators, \\p{..}, and a pattern whose DFA
    # is too large for a device table (the device reports every key, the host decides with the NFA)
    ("p2pkh", r"at\b", False), ("p2pkh", r"^\b1C\B", False), ("p2wpkh", "(?x) q q $ # two q", False), ("p2pkh", "[a-z&&[^aeiou]]{4}$", False),
    ("ethereum", r"\p{Lu}{4}$", False), ("p2sh-p2wpkh", "[A-Z--[A-M]]{3}", False), ("p2pkh", "a.{13}$", False, 1 << 14),
    # a prefix and a tail: interval prefilter + device DFA on the survivors (vg_matcher_dev::confirm_dfa)
    ("p2pkh", "^1C.*z$", False), ("p2wpkh", "^bc1qa.*zz$", False), ("ethereum", "^0xAb", False), ("p2sh-p2wpkh", "^3A.*[0-9]$", False),
]


@pytest.mark.parametrize("case", SMALL_CASES, ids=lambda c: "%s-%s%s" % (c[0], c[1], "-i" if c[2] else ""))
def test_scan_match_sets_small(small_runner, case):
    fmt, pat, ci = case[:3]
    n = case[3] if len(case) > 3 else 1 << 18
    res = _check_scan(small_runner, fmt, pat, 2**64, n, ci)
    assert len(res.matches) > 0, "vacuous case: resize it"


@pytest.mark.parametrize("fmt,pat,kind", [("p2tr", "^bc1q", "none"), ("p2pkh", "^3", "none"), ("p2wpkh", "^1", "none"), ("ethereum", "^0xg", "none"),
                                          ("p2sh-p2wpkh", "^(3Ab|3Cd)$", "intervals")])
def test_scan_patterns_that_cannot_match(small_runner, fmt, pat, kind):
    """Patterns no address of the format can match. Where the lowering can tell (a prefix outside the format's head or alphabet) it
    says "none" and the scan reports the range as scanned without a launch; where it cannot (a 3-character exact string: every
    address is longer) the interval prefilter + device DFA find nothing. Either way: no matches, every key counted."""
    res = _check_scan(small_runner, fmt, pat, 2**64, 1 << 16)
    assert res.matches == [] and res.operations == 1 << 16
    assert small_runner.matcher_kind() == kind


def test_scan_edges(small_runner):
    # range that includes scalar 0 (skipped, not counted) and dense pattern `.` (range's default, src/lib.rs:519)
    res = _scan(small_runner, "p2pkh", ".", 0, 3000)
    assert [int(m.hex, 16) for m in res.matches] == list(range(1, 3000)) and res.operations == 2999
    # tail of the group: stops at n-1
    cfg = vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=vg.UNLIMITED, start=N - 500, end=None)
    res = vg.scan_gpu_with_runner(vg.Pattern("^1"), cfg, None, None, small_runner)
    assert [int(m.hex, 16) for m in res.matches] == list(range(N - 500, N)) and res.operations == 500
    # count: first `count` matches in ascending key order (src/gpu.rs:1095-1104)
    res = _scan(small_runner, "p2pkh", "^1A", 1, 1 << 16, count=3)
    keys, _, _ = oracle.scan_range("p2pkh", "^1A", 1, 1 << 16)
    assert [int(m.hex, 16) for m in res.matches] == keys[:3]
    # count == 0: the library stops before the first window (src/gpu.rs:981-996); only the CLI maps 0 to "no limit" (src/lib.rs:524)
    res = _scan(small_runner, "p2pkh", "^1A", 1, 1 << 16, count=0)
    assert list(res.matches) == [] and res.operations == 0
    # single key range, ragged end
    res = _scan(small_runner, "p2pkh", ".", 77, 1)
    assert [int(m.hex, 16) for m in res.matches] == [77]
    # end < start is rejected cleanly (the reference underflows, src/scanner.rs:228)
    with pytest.raises(Exception):
        small_runner.scan_raw(10, 5)


def test_scan_exact_address_planted(runner):
    # C5 shape at reduced size: exact-address pattern for a planted key inside a multi-window range
    start = 2**65
    n = runner.batch_size * 2 + 12345
    off = int(0.83 * n)
    addr = oracle.address("p2pkh", start + off)[0]
    res = _scan(runner, "p2pkh", "^" + addr + "$", start, n)
    assert [int(m.hex, 16) for m in res.matches] == [start + off]
    assert res.matches[0].address == addr and res.operations == n
    assert runner.matcher_kind() == "exact"


def test_p2tr_published_bip341_vector(small_runner, published):
    """BIP341 wallet-test-vectors: internalPrivkey -> output key -> bc1p address, through both GPU stages (walk + TapTweak)."""
    v = published["taproot"]["vectors"][0]
    k = int(v["internal_privkey"], 16)
    assert small_runner.digest_batch("p2tr", k, 1).hex() == v["output_key"]
    res = _scan(small_runner, "p2tr", "^" + v["address"] + "$", k - 1000, 3000)
    assert [(int(m.hex, 16), m.address) for m in res.matches] == [(k, v["address"])]
    small_runner.dispatch_p2tr(k, 0, n=8)
    assert small_runner.await_result_p2tr(0)[0][0].hex() == v["internal_pubkey"]


def test_published_key_to_address_vectors(small_runner, published):
    """Bitcoin wiki worked example (private key -> compressed public key -> P2PKH address) and the WIF example, on the GPU path."""
    v = published["base58check_p2pkh"]["vectors"][1]
    k = int(v["key"], 16)
    assert small_runner.generate_batch(k, 1)[0].hex() == v["hash160"]
    res = _scan(small_runner, "p2pkh", "^" + v["address"] + "$", k - 5, 10)
    assert [(m.address, int(m.hex, 16)) for m in res.matches] == [(v["address"], k)]
    w = published["wif"]["vectors"][0]
    res = _scan(small_runner, "p2pkh-uncompressed", ".", int(w["key"], 16), 1)
    assert res.matches[0].wif == w["wif_uncompressed"] and res.matches[0].address == "1GAehh7TsJAHuUAeKZcXf5CnwuGuGgyX2S"
    res = _scan(small_runner, "p2pkh", ".", int(w["key"], 16), 1)
    assert res.matches[0].wif == w["wif_compressed"]


def test_p2tr_exact_address_planted(small_runner):
    start = 2**65
    n = small_runner.batch_size * 2 + 345
    off = int(0.61 * n)
    addr = oracle.address("p2tr", start + off)[0]
    res = _scan(small_runner, "p2tr", "^" + addr + "$", start, n)
    assert [int(m.hex, 16) for m in res.matches] == [start + off] and res.matches[0].address == addr and res.operations == n
    assert small_runner.matcher_kind() == "exact"


def test_p2tr_long_job_builds_the_wide_table(runner):
    """A P2TR job of at least 2^30 keys switches the tweak stage from the 15-window table (67 MB) to the 10-window one of signed
    digits (6 x 26 + 4 x 25 bits, 17 GB in HBM; engine.cpp ensure_taproot). Same results as the oracle, through the scan and the
    digest path."""
    MID = 0x40000000 | (6 << 8) | 25
    import os
    start, n = 2**70, (1 << 32) + 12345
    cfg = vg.ScanConfig(format=vg.AddressFormat.P2tr, count=5, start=start, end=start + n - 1)
    res = vg.scan_gpu_with_runner(vg.Pattern("^bc1pqq"), cfg, None, None, runner)
    if "VGEN_TAP_BITS" not in os.environ:
        assert runner.tap_bits() == MID
    keys, _, _ = oracle.scan_range("p2tr", "^bc1pqq", start, 30000)
    assert len(keys) >= 5 and [int(m.hex, 16) for m in res.matches] == keys[:5]
    for m in res.matches:
        assert m.address == oracle.address("p2tr", int(m.hex, 16))[0]
    got = runner.digest_batch("p2tr", start + 777, 3000)          # the wide table stays in place for short calls
    assert got == oracle.digest_batch("p2tr", start + 777, 3000)
    assert "VGEN_TAP_BITS" in os.environ or runner.tap_bits() == MID


@pytest.mark.parametrize("geom", ["s7.4", "s5.1", "s4.0", "13"])
def test_p2tr_table_geometries(monkeypatch, geom):
    """The tweak stage's fixed-base table in other geometries (scan_types.h; VGEN_TAP_BITS): signed mixed-radix windows — the shape
    of the 111.7 GB 9-window table for the longest jobs, here with 36, 51 and 64 small windows — and an odd plain width. Witness
    programs equal to the oracle's at both ends of the scalar range; a scan finds the oracle's matches."""
    monkeypatch.setenv("VGEN_TAP_BITS", geom)
    r = vg.GpuRunner(vg.DEFAULT_BATCH_SIZE)
    try:
        for start, n in ((1, 3000), (2**200 + 12345, 3000), (N - 2000, 3000)):
            got = r.digest_batch("p2tr", start, n)
            assert got == oracle.digest_batch("p2tr", start, n), (geom, hex(start))
        want_bits = (0x40000000 | (int(geom[1:].split(".")[1]) << 8) | int(geom[1:].split(".")[0])) if geom[0] == "s" else int(geom)
        assert r.tap_bits() == want_bits
        _check_scan(r, "p2tr", "^bc1pq", 2**64, 1 << 14)
    finally:
        r.close()


def test_config_c1_full_match_set(runner):
    """BASELINE.json config C1: 2^24-key window from 2^64, P2PKH, ^1Cat — full match set vs the oracle."""
    res = _check_scan(runner, "p2pkh", "^1Cat", 2**64, 1 << 24)
    assert 150 < len(res.matches) < 300          # ~217 expected
    assert all(m.address.startswith("1Cat") for m in res.matches)


def _verify_hits(fmt, pat, res, start, n, ci=False):
    """Soundness of a scan too large for a full oracle pass: ascending unique keys inside the range, and every reported hit
    re-derived by the oracle (address, WIF) and matched by the oracle's regex."""
    keys = [int(m.hex, 16) for m in res.matches]
    assert keys == sorted(set(keys)) and all(start <= k < start + n for k in keys)
    assert res.operations == n
    for m in res.matches:
        a, w, _ = oracle.address(fmt, int(m.hex, 16))
        assert (a, w) == (m.address, m.wif) and oracle.regex_match(pat, a, ci), m
    return keys


def test_config_c2_catdog_planted(runner):
    """BASELINE.json config C2: P2PKH ^1CatDog (hash160 interval path) over 2^34 keys from 2^64. The pattern itself expects
    ~1 hit, so completeness is shown with planted keys: the 8-character prefixes of three keys of the range, as an alternation
    (lowered to intervals from the DFA), must each be found; every reported hit is re-derived by the oracle; the
    oracle-scanned head of the range is equal."""
    S, n = 2**64, 1 << 34
    res = _scan(runner, "p2pkh", "^1CatDog", S, n)
    assert runner.matcher_kind() == "intervals"
    keys = _verify_hits("p2pkh", "^1CatDog", res, S, n)
    want, _, _ = oracle.scan_range("p2pkh", "^1CatDog", S, 1 << 20)
    assert [k for k in keys if k < S + (1 << 20)] == want
    planted = [S + 12345, S + n // 3, S + n - 7]
    pat = "^(" + "|".join(oracle.address("p2pkh", k)[0][:8] for k in planted) + ")"
    res = _scan(runner, "p2pkh", pat, S, n)
    assert runner.matcher_kind() == "intervals"
    keys = _verify_hits("p2pkh", pat, res, S, n)
    assert all(k in keys for k in planted), "planted keys missing"


def test_config_c3_p2wpkh_dead_suffix_full_match_set(runner):
    """BASELINE.json config C3: P2WPKH dead$ (on-GPU Bech32 + shared-memory DFA) — full match set over 2^24 keys from 2^64
    against the oracle (probability 2^-20 per key: ~16 hits, so the comparison is not vacuous)."""
    res = _check_scan(runner, "p2wpkh", "dead$", 2**64, 1 << 24)
    assert len(res.matches) > 0 and all(m.address.endswith("dead") for m in res.matches)
    assert runner.matcher_kind() == "dfa"


def test_config_c4_ethereum_dead_prefix_full_match_set(runner):
    """BASELINE.json config C4: Ethereum ^0xdead (uncompressed key + Keccak-256; the EIP-55 case of the four letters must come
    out lower-case, a 1/16 sub-filter) — full match set over 2^24 keys from 2^64 against the oracle."""
    res = _check_scan(runner, "ethereum", "^0xdead", 2**64, 1 << 24)
    assert len(res.matches) > 0 and all(m.address.startswith("0xdead") for m in res.matches)


def test_config_c5_exact_address_planted(runner):
    """BASELINE.json config C5 shape: exact-address match over a 2^35-key window from 2^65, the planted key at 0.83 of the
    window: exactly that key comes back and every key of the window is counted (the 2^40 window: test_full_size_configs)."""
    start, n = 2**65, 1 << 35
    off = int(0.83 * n)
    addr = oracle.address("p2pkh", start + off)[0]
    res = _scan(runner, "p2pkh", "^" + addr + "$", start, n)
    assert [int(m.hex, 16) for m in res.matches] == [start + off] and res.matches[0].address == addr
    assert res.operations == n and runner.matcher_kind() == "exact"


def test_sharded_scan_equals_single(runner):
    """Two shard contexts (as two torchrun ranks would hold) cover the range disjointly: union == single-context set."""
    start, n = 2**64, runner.batch_size * 4 + 999
    single = _scan(runner, "p2pkh", "^1Ca", start, n)
    parts = []
    ops = 0
    for i in range(2):
        r = vg.GpuRunner(shard=(i, 2))
        try:
            res = _scan(r, "p2pkh", "^1Ca", start, n)
            parts += [int(m.hex, 16) for m in res.matches]
            ops += res.operations
        finally:
            r.close()
    assert sorted(parts) == [int(m.hex, 16) for m in single.matches]
    assert len(set(parts)) == len(parts) and ops == n


def test_stop_flag_and_progress(small_runner):
    stop = vg.StopFlag()
    seen = []

    def cb(k):
        seen.append(k)
        if len(seen) == 3:
            stop.set()
    cfg = vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=vg.UNLIMITED, start=2**64, end=2**64 + (1 << 30))
    res = vg.scan_gpu_with_runner(vg.Pattern("^1CatDogs"), cfg, cb, stop, small_runner)
    assert seen == sorted(seen) and len(seen) >= 3
    assert res.operations < (1 << 30)


def _n_gpus():
    from vgen_b200 import _cabi
    return int(_cabi.lib().vgen_cuda_device_count())


# device lists of one context. [0, 0] drives the same GPU as two "devices" (two host threads, four lanes): the multi-device
# machinery — one host thread per device, hits merged in window order on the calling thread — runs on a 1-GPU box too.
MULTI = [[0, 0], [0, 1], [0, 1, 2, 3], [0, 1, 2, 3, 4, 5, 6, 7]]


@pytest.mark.parametrize("devs", MULTI, ids=lambda d: "x".join(map(str, d)))
def test_multi_device_context_equals_single(runner, devs):
    """One context over several GPUs (one host thread per device, hits merged on the host) == the single-GPU set:
    the path `vgen --devices 0,1,...` and the Rust binding use."""
    if max(devs) >= _n_gpus():
        pytest.skip("needs %d GPUs" % (max(devs) + 1))
    start, n = 2**64, runner.batch_size * (2 * len(devs) + 3) + 4321
    single = _scan(runner, "p2pkh", "^1Ca", start, n)
    rm = vg.GpuRunner(devices=devs)
    try:
        assert rm.lane_count() == 2 * len(devs) and rm.stats()["devices"] == len(devs)
        multi = _scan(rm, "p2pkh", "^1Ca", start, n)
        assert [m.hex for m in multi.matches] == [m.hex for m in single.matches] and multi.operations == n
        # a hit limit: the first `count` matches in ascending key order, whichever device found them
        few = _scan(rm, "p2pkh", "^1Ca", start, n, count=7)
        assert [m.hex for m in few.matches] == [m.hex for m in single.matches][:7]
        # a DFA pattern and a dense one through the same context
        _check_scan(rm, "p2wpkh", "qq$", 2**64, 1 << 18)
        res = _scan(rm, "p2pkh", ".", 5, 5000)
        assert [int(m.hex, 16) for m in res.matches] == list(range(5, 5005))
        # P2TR: stage 2 derives every key's scalar from its window's base, whichever lane runs the window
        _check_scan(rm, "p2tr", "^bc1pq", 2**64 + 99, 1 << 15)
    finally:
        rm.close()


def test_failed_device_windows_are_requeued(runner, monkeypatch):
    """SURVEY.md section 5, failure row: a device whose CUDA call fails drops out of the scan, the windows it had in flight
    and the rest of its lanes' windows are taken over by the surviving devices, and the result is still the exact set
    (the reference falls back to the CPU scanner instead, src/lib.rs:727-746,1185-1197; north_star forbids that).
    VGEN_FAULT_INJECT=<device index>:<n> makes the n-th window issue of that device return an error."""
    devs = [0, 1] if _n_gpus() >= 2 else [0, 0]
    start, n = 2**64, runner.batch_size * 9 + 777
    single = _scan(runner, "p2pkh", "^1Ca", start, n)
    rm = vg.GpuRunner(devices=devs)
    try:
        monkeypatch.setenv("VGEN_FAULT_INJECT", "1:2")          # device 1 fails at its third window
        res = _scan(rm, "p2pkh", "^1Ca", start, n)
        st = rm.stats()
        assert st["failed_devices"] == 1 and st["requeued_windows"] >= 1, st
        assert [m.hex for m in res.matches] == [m.hex for m in single.matches] and res.operations == n
        monkeypatch.delenv("VGEN_FAULT_INJECT")
        # the device stays out; later scans of the context run on the survivor alone and stay exact
        res = _scan(rm, "p2pkh", "^1Ca", start + 99, n)
        keys, _, _ = oracle.scan_range("p2pkh", "^1Ca", start + 99, 1 << 20)
        assert [int(m.hex, 16) for m in res.matches if int(m.hex, 16) < start + 99 + (1 << 20)] == keys and res.operations == n
        assert rm.stats()["failed_devices"] == 1
        # when the last device fails too, the scan fails loudly: there is no CPU fallback
        monkeypatch.setenv("VGEN_FAULT_INJECT", "0:1")
        with pytest.raises(Exception, match="no device left"):
            _scan(rm, "p2pkh", "^1Ca", start, n)
    finally:
        rm.close()


def test_hit_buffer_overflow_is_rescanned(monkeypatch):
    """Hit buffers hold fewer records than a window has keys; a window that reports more hits than fit is re-run in pieces of a
    smaller ramp level (engine.cpp harvest). Forced here with a tiny buffer (VGEN_HIT_CAP) and a pattern that matches ~15 % of
    the keys; the match set must still equal the oracle's."""
    monkeypatch.setenv("VGEN_HIT_CAP", str(1 << 18))
    r = vg.GpuRunner(8 * 262144)                       # 2,097,152 keys per window, 262,144 hit records per slot
    try:
        assert r.stats()["hit_cap"] == 1 << 18 and r.batch_size == 8 * 262144
        n = 2 * r.batch_size + 1000
        res = _check_scan(r, "p2pkh", "[1-9]$", 2**64, n)
        assert len(res.matches) > 2 * (1 << 18)
        assert r.stats()["overflow_rescans"] >= 2
    finally:
        r.close()


def test_big_text_records_are_copied_out_whole(small_runner):
    """Tens of thousands of hits: the records reach the caller through a bytes object the library fills on several threads
    (vgen_cuda_scan_fetch_text with the exact size; gpu.py fetch_text). Same bytes as the library's own blob, every record right."""
    start, n = 2**64 + 777, 1 << 15
    res = _scan(small_runner, "p2pkh", "^1", start, n)                       # every P2PKH address starts with 1
    assert len(res.matches) == n and res.operations == n
    ptr, ln = C.c_void_p(), C.c_uint64(0)
    L = small_runner._lib
    assert L.vgen_cuda_scan_text(small_runner._ctx, C.cast(C.byref(ptr), C.POINTER(C.c_char_p)), C.byref(ln)) == 1
    blob = C.string_at(ptr, ln.value)
    assert ln.value > (4 << 20)                                              # large enough for the threaded copy
    text, got = small_runner.fetch_text(n)
    assert got == n and bytes(text) == blob
    for i in (0, 1, 4097, n // 2, n - 2, n - 1):
        m = res.matches[i]
        a, w, _ = oracle.address("p2pkh", start + i)
        assert (int(m.hex, 16), m.address, m.wif) == (start + i, a, w)


def test_two_lanes_per_gpu_cover_the_range_once(runner):
    """Every GPU is driven through 2 lanes (streams) taking alternate windows: the union is the range, each key once."""
    assert runner.lane_count() == 2 * len(runner.devices)
    start, n = 2**64 + 12345, runner.batch_size * 5 + 77          # odd number of windows + a ragged tail
    res = _scan(runner, "p2pkh", "^1Ab", start, n)
    keys, ops, _ = oracle.scan_range("p2pkh", "^1Ab", start, 1 << 22)
    got = [int(m.hex, 16) for m in res.matches]
    assert got == sorted(set(got)) and res.operations == n
    assert [k for k in got if k < start + (1 << 22)] == keys      # the oracle-checkable head of the range
    # every reported hit re-derived by the oracle (soundness over the full range)
    for m in res.matches[:: max(1, len(res.matches) // 200)]:
        assert oracle.address("p2pkh", int(m.hex, 16))[0] == m.address and m.address.startswith("1Ab")


def test_dense_prefix_with_hit_limit_uses_small_windows(runner):
    """generate -p ^1A -c 1 (src/lib.rs:347): a dense interval matcher with a hit limit starts on the ramp levels, so the
    first match comes back after a small window instead of a full-size round of 2 x 77.6 M keys."""
    cfg = vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=1, start=2**64 + 5, end=None)
    res = vg.scan_gpu_with_runner(vg.Pattern("^1A"), cfg, None, None, runner)
    keys, _, _ = oracle.scan_range("p2pkh", "^1A", 2**64 + 5, 1 << 12)
    assert [int(m.hex, 16) for m in res.matches] == keys[:1]
    assert res.operations < runner.batch_size
    # a sparse prefix with the same limit goes straight to full-size windows (no ramp): operations are whole windows
    cfg = vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=1, start=2**64, end=2**64 + 3 * runner.batch_size - 1)
    res = vg.scan_gpu_with_runner(vg.Pattern("^1CatDogCat"), cfg, None, None, runner)
    assert res.matches == [] and res.operations == 3 * runner.batch_size


def test_dispatch_await_frames(small_runner):
    """The reference's two-frame API (src/gpu.rs:535-658, 667-807): frames dispatched back to back, awaited in any order."""
    r = small_runner
    a, b = 2**64 + 7, 2**64 + 7 + r.batch_size
    r.dispatch(a, 0)
    r.dispatch(b, 1)
    h1, s1 = r.await_result(1)
    h0, s0 = r.await_result(0)
    assert s0 == a.to_bytes(32, "big") and s1 == b.to_bytes(32, "big") and len(h0) == r.batch_size
    assert b"".join(h0[:3000]) == oracle.hash160_batch(a, 3000) and b"".join(h1[-3000:]) == oracle.hash160_batch(b + r.batch_size - 3000, 3000)
    # P2TR frame: the UNTWEAKED even-Y internal key = x of the public key
    r.dispatch_p2tr(a, 0)
    xs, _ = r.await_result_p2tr(0)
    for i in (0, 1, 2, 1000, r.batch_size - 1):
        assert int.from_bytes(xs[i], "big") == oracle.pubkey(a + i)[0]
    res = vg.scan_gpu_p2tr_with_runner(vg.Pattern("^bc1pqq"), vg.ScanConfig(format=vg.AddressFormat.P2pkh, count=vg.UNLIMITED, start=a, end=a + (1 << 16)), None, None, r)
    keys, _, _ = oracle.scan_range("p2tr", "^bc1pqq", a, (1 << 16) + 1)
    assert [int(m.hex, 16) for m in res.matches] == keys and all(m.format == vg.AddressFormat.P2tr for m in res.matches)
