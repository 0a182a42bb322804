"""vgen_cuda_hash160_batch_static (include/vgen_cuda.h): the allocation-free hash160 batch, whose kernels work in a file-scope
__device__ array of the library instead of cudaMalloc'ed buffers. Called through the C ABI and compared, bit for bit, with the
existing entry point vgen_cuda_hash160_batch on several calls in a row (the array keeps the previous call's contents).

Collected last on purpose (file name)."""
import ctypes as C
import random

import pytest

import oracle
from vgen_b200 import _cabi

pytestmark = pytest.mark.gpu
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
MAX_KEYS = 1 << 20          # VGEN_STATIC_MAX_KEYS


def _be(k):
    return int(k).to_bytes(32, "big")


@pytest.fixture(scope="module")
def ctx():
    L = _cabi.lib()
    h = C.c_void_p()
    assert L.vgen_cuda_create(None, 0, 0, C.byref(h)) == 0, L.vgen_cuda_last_error(None)
    yield h
    L.vgen_cuda_destroy(h)


def _existing(ctx, start, n):
    L = _cabi.lib()
    buf = C.create_string_buffer(n * 20)
    assert L.vgen_cuda_hash160_batch(ctx, _be(start), n, buf) == 0, L.vgen_cuda_last_error(ctx)
    return buf.raw[:n * 20]


def _static(start, n, stream=None, poison=0xEE):
    """one call of the new entry point into caller-owned device memory (a torch tensor, filled with `poison` first)"""
    import torch
    L = _cabi.lib()
    out = torch.full((max(n, 1) * 20,), poison, dtype=torch.uint8, device="cuda:0")
    err = C.create_string_buffer(256)
    if stream is None:
        rc = L.vgen_cuda_hash160_batch_static(0, _be(start), n, out.data_ptr(), None, err, len(err))
    else:
        stream.wait_stream(torch.cuda.current_stream())          # the fill above ran on torch's current stream
        rc = L.vgen_cuda_hash160_batch_static(0, _be(start), n, out.data_ptr(), stream.cuda_stream, err, len(err))
        stream.synchronize()
    assert rc == 0, err.value
    return bytes(out.cpu().numpy().tobytes()[:n * 20])


def test_static_batch_matches_existing_entry_point_on_calls_in_a_row(ctx):
    import torch
    rnd = random.Random(20261105)
    a = rnd.getrandbits(255) | 1
    # 1st and 2nd call: different starts, the second smaller than the first (stale scratch / centres / counters must not show)
    calls = [(a, 70_000), (rnd.getrandbits(200), 33_333),
             (0, 5_000),                 # scalar 0: zeros, and its batch goes through the slow pass
             (N - 3_000, 5_000),         # the last 3,000 valid scalars, then zeros
             (N + 5, 100),               # nothing valid
             (a, 70_000)]                # the first call again, after the others
    got = []
    for start, n in calls:
        g = _static(start, n)
        assert g == _existing(ctx, start, n), (hex(start), n)
        got.append(g)
    assert got[0] == got[-1]
    assert got[2][:20] == b"\0" * 20 and got[3][3_000 * 20:] == b"\0" * 40_000 and got[4] == b"\0" * 2_000
    # anchored to the oracle as well (a prefix: the oracle is a scalar CPU port)
    assert got[0][:20 * 2_048] == oracle.hash160_batch(a, 2_048)
    # on a caller-owned stream, twice in a row
    s = torch.cuda.Stream(device="cuda:0")
    for start, n in [(a + 12_345, 9_000), (a + 999, 1_025)]:
        assert _static(start, n, stream=s) == _existing(ctx, start, n)


def test_static_batch_full_size_and_limit(ctx):
    L = _cabi.lib()
    start = 0x2832ED74F2B5E35EE - 1000
    full = _static(start, MAX_KEYS)
    assert full == _existing(ctx, start, MAX_KEYS)
    again = _static(start + 1, MAX_KEYS)                 # second full-size call in a row, shifted by one key
    assert again[:-20] == full[20:]
    # beyond the maximum: a clear error, nothing written
    import torch
    out = torch.full((64,), 0xEE, dtype=torch.uint8, device="cuda:0")
    err = C.create_string_buffer(256)
    rc = L.vgen_cuda_hash160_batch_static(0, _be(1), MAX_KEYS + 1, out.data_ptr(), None, err, len(err))
    assert rc == -1 and b"VGEN_STATIC_MAX_KEYS" in err.value and str(MAX_KEYS + 1).encode() in err.value
    assert bytes(out.cpu().numpy().tobytes()) == b"\xee" * 64
    # bad arguments
    assert L.vgen_cuda_hash160_batch_static(99, _be(1), 16, out.data_ptr(), None, err, len(err)) == -1 and b"device" in err.value
    assert L.vgen_cuda_hash160_batch_static(0, _be(1), 16, None, None, err, len(err)) == -1
    assert L.vgen_cuda_hash160_batch_static(0, _be(1), 0, None, None, err, len(err)) == 0
