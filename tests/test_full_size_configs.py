"""BASELINE.json configs C2-C5 at their full single-GPU sizes (2^36 / 2^34 / 2^34 / 2^40 keys), checked through
size-independent properties: soundness (every reported hit is re-derived by the oracle and matches the pattern), completeness
on planted keys and on an oracle-scanned head of the range, exact operation counts.

Collected by pytest (`-m gpu`; also marked `slow`: about four minutes on one B200, most of it C5's 2^40 keys). With several GPUs
visible the scans run in ONE context over all of them (hits merged on the host), which is the path `vgen --devices 0,1,...` uses.
Also runnable by hand:  python tests/test_full_size_configs.py [c2 c3 c4 c5]
"""
import os
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import oracle          # noqa: E402  (tests/ may use the oracle as the checker)
import vgen_b200 as vg  # noqa: E402

pytestmark = [pytest.mark.gpu, pytest.mark.slow]
S = 1 << 64


@pytest.fixture(scope="module")
def runner():
    from vgen_b200 import _cabi
    n = max(1, int(_cabi.lib().vgen_cuda_device_count()))
    r = vg.GpuRunner(devices=list(range(n)))
    yield r
    r.close()


def scan(r, fmt, pat, start, n):
    t0 = time.perf_counter()
    res = vg.scan_gpu_with_runner(vg.Pattern(pat), vg.ScanConfig(format=vg.AddressFormat.parse(fmt), count=vg.UNLIMITED, start=start, end=start + n - 1), None, None, r)
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


def test_c2_p2pkh_catdog_2_36(runner):
    n = 1 << 36
    planted = [S + 12345, S + n // 3, S + n - 7]
    res, dt = scan(runner, "p2pkh", "^1CatDog", S, n)
    nh, nw = check_hits("p2pkh", "^1CatDog", res, S, n, 1 << 22)
    print("C2 p2pkh ^1CatDog 2^36 keys: %d hits (all verified), %.1f s, %.0f Mkeys/s" % (nh, dt, n / dt / 1e6), flush=True)
    pat = "^(" + "|".join(oracle.address("p2pkh", k)[0][:8] for k in planted) + ")"
    res, dt = scan(runner, "p2pkh", pat, S, n)
    keys = [int(m.hex, 16) for m in res.matches]
    assert all(k in keys for k in planted), "planted keys missing"
    check_hits("p2pkh", pat, res, S, n, 1 << 20)
    print("C2 planted 8-char prefixes of 3 keys: found all, %d hits, matcher %s, %.1f s" % (len(keys), runner.matcher_kind(), dt), flush=True)


def test_c3_p2wpkh_dead_suffix_2_34(runner):
    n = 1 << 34
    res, dt = scan(runner, "p2wpkh", "dead$", S, n)
    nh, nw = check_hits("p2wpkh", "dead$", res, S, n, 1 << 23)
    assert nh > 10000 and nw > 0          # ~16,384 expected
    print("C3 p2wpkh dead$ 2^34 keys: %d hits (all verified; %d in the oracle-scanned head), %.1f s, %.0f Mkeys/s" % (nh, nw, dt, n / dt / 1e6), flush=True)


def test_c4_ethereum_dead_prefix_2_34(runner):
    n = 1 << 34
    res, dt = scan(runner, "ethereum", "^0xdead", S, n)
    nh, nw = check_hits("ethereum", "^0xdead", res, S, n, 1 << 22)
    assert nh > 10000          # 2^34 / 16^4 / 16 = 16,384 expected
    print("C4 ethereum ^0xdead 2^34 keys: %d hits (all verified, EIP-55 lower-case only; %d in the head), %.1f s, %.0f Mkeys/s" % (nh, nw, dt, n / dt / 1e6), flush=True)


def test_c5_exact_address_2_40(runner):
    start, n = 1 << 65, 1 << 40
    off = int(0.83 * n)
    addr = oracle.address("p2pkh", start + off)[0]
    res, dt = scan(runner, "p2pkh", "^" + addr + "$", start, n)
    assert [int(m.hex, 16) for m in res.matches] == [start + off] and res.matches[0].address == addr
    print("C5 exact address over 2^40 keys from 2^65 on %d GPU(s): the planted key at offset %d found, nothing else, %.1f s, %.0f Mkeys/s (matcher %s)"
          % (len(runner.devices), off, dt, n / dt / 1e6, runner.matcher_kind()), flush=True)


if __name__ == "__main__":
    which = [a.lower() for a in sys.argv[1:]] or ["c2", "c3", "c4", "c5"]
    from vgen_b200 import _cabi
    r = vg.GpuRunner(devices=list(range(max(1, int(_cabi.lib().vgen_cuda_device_count())))))
    for name, fn in (("c2", test_c2_p2pkh_catdog_2_36), ("c3", test_c3_p2wpkh_dead_suffix_2_34), ("c4", test_c4_ethereum_dead_prefix_2_34), ("c5", test_c5_exact_address_2_40)):
        if name in which:
            fn(r)
    r.close()
