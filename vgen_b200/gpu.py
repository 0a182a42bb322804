"""GpuRunner / scan_gpu_with_runner / scan_gpu — the reference's GPU scan surface (src/gpu.rs:18-63,128-136,138,
535-665,920-1166) over libvgen_cuda.so. Same names and argument meaning; the wgpu backends are replaced by CUDA, and
there is no CPU fallback: if the library or a B200 is missing, construction raises."""
import collections.abc
import ctypes as C
import enum
import itertools
import secrets
import threading
import time

from . import _cabi
from .address import AddressFormat, GeneratedAddress, address_from_digest, wif_from_key
from .pattern import Pattern
from .scanner import MatchList, ScanConfig, ScanResult

DEFAULT_BATCH_SIZE = 512 * 1024          # src/gpu.rs:83; here only a keys-per-launch hint (0 = fill the GPU)
UNLIMITED = (1 << 64) - 1                # usize::MAX: what the reference's CLI turns `--count 0` into (src/lib.rs:524)
SECP256K1_N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141


class GpuBackend(enum.Enum):             # src/gpu.rs:18-26, plus Cuda; the graphics backends do not exist here
    Auto = "auto"
    Cuda = "CUDA"
    Vulkan = "Vulkan"
    Metal = "Metal"
    Dx12 = "DX12"
    Gl = "OpenGL"

    def name_str(self):
        return self.value

    def __str__(self):
        return self.value


def _be(k):
    return int(k).to_bytes(32, "big")


class GpuRunner:
    """GpuRunner::new(batch_size, backend) (src/gpu.rs:138). `devices` selects CUDA devices (default [0]);
    `shard=(index, count)` lets one process per GPU split a range with disjoint start scalars and a common stride."""

    def __init__(self, batch_size=0, backend=GpuBackend.Auto, devices=None, shard=None, stream=None):
        if backend not in (GpuBackend.Auto, GpuBackend.Cuda):
            raise RuntimeError("backend %s is not available: this build drives B200 GPUs through CUDA only" % backend)
        L = _cabi.lib()
        devs = list(devices) if devices else [0]
        arr = (C.c_int * len(devs))(*devs)
        ctx = C.c_void_p()
        rc = L.vgen_cuda_create(arr, len(devs), int(batch_size or 0), C.byref(ctx))
        if rc != 0:
            raise _cabi.VgenError(rc, (L.vgen_cuda_last_error(None) or b"").decode())
        self._ctx = ctx
        self._lib = L
        self.devices = devs
        self.batch_size = int(L.vgen_cuda_window_keys(ctx))
        name = C.create_string_buffer(256)
        L.vgen_cuda_device_name(devs[0], name, 256)
        self.device_name = name.value.decode()
        self._backend = GpuBackend.Cuda
        if shard is not None:
            self._check(L.vgen_cuda_set_shard(ctx, int(shard[0]), int(shard[1])))
        if stream is not None:
            self._check(L.vgen_cuda_set_stream(ctx, 0, C.c_void_p(int(stream))))

    @staticmethod
    def new(batch_size=0, backend=GpuBackend.Auto, **kw):
        return GpuRunner(batch_size, backend, **kw)

    def backend(self):                                    # src/gpu.rs:134-136
        return self._backend

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.vgen_cuda_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise _cabi.VgenError(rc, (self._lib.vgen_cuda_last_error(self._ctx) or b"").decode())

    # -- reference-shaped batch API (generate_batch, src/gpu.rs:660-665): hash160 of n sequential keys.
    # The frame size of this API is independent of the scan window (which defaults to 77.6 M keys here): `frame_size` keys,
    # DEFAULT_BATCH_SIZE unless the runner was created with a smaller batch. Results come back as ONE buffer (DigestBatch).
    @property
    def frame_size(self):
        return min(self.batch_size, DEFAULT_BATCH_SIZE)

    def generate_batch(self, start_key, n=None):
        n = self.frame_size if n is None else int(n)
        buf = C.create_string_buffer(20 * n)
        self._check(self._lib.vgen_cuda_hash160_batch(self._ctx, _be(start_key), n, buf))
        return DigestBatch(buf.raw, 20)

    # -- the reference's two-frame dispatch / await API (src/gpu.rs:535-658, 667-807). A frame is computed by a worker thread
    # (ctypes releases the GIL); frames of one runner are serialised by a lock because a context runs one job at a time.
    def _dispatch(self, kind, start_key, frame_index, n=None):
        if frame_index not in (0, 1):
            raise IndexError("frame_index must be 0 or 1")          # two frames in flight (src/gpu.rs:399)
        if not hasattr(self, "_frames"):
            self._frames, self._frame_lock = {}, threading.Lock()
        n = self.frame_size if n is None else int(n)
        box = {}

        def work():
            try:
                with self._frame_lock:
                    if kind == "p2tr":
                        buf = C.create_string_buffer(32 * n)
                        self._check(self._lib.vgen_cuda_digest_batch(self._ctx, 6, _be(start_key), n, buf, 32))
                        box["out"] = DigestBatch(buf.raw, 32)
                    else:
                        box["out"] = self.generate_batch(start_key, n)
            except Exception as e:                                   # surfaced by await_result
                box["err"] = e
        t = threading.Thread(target=work, daemon=True)
        t.start()
        self._frames[(kind, frame_index)] = (t, box, _be(start_key))

    def _await(self, kind, frame_index):
        t, box, start = self._frames.pop((kind, frame_index))
        t.join()
        if "err" in box:
            raise box["err"]
        return box["out"], start

    def dispatch(self, start_key, frame_index=0, n=None):
        self._dispatch("h160", start_key, frame_index, n)

    def await_result(self, frame_index=0):
        """(DigestBatch of `frame_size` hash160s, start_key_bytes) of the frame (src/gpu.rs:602-658)."""
        return self._await("h160", frame_index)

    dispatch_batch_hash = dispatch                                   # src/gpu.rs:809: same result through the batch-inversion kernel

    def dispatch_p2tr(self, start_key, frame_index=0, n=None):
        self._dispatch("p2tr", start_key, frame_index, n)

    def await_result_p2tr(self, frame_index=0):
        """(DigestBatch of 32-byte x-only even-Y internal keys, start_key_bytes): untweaked, as src/gpu.rs:750-807."""
        return self._await("p2tr", frame_index)

    def digest_batch(self, fmt, start_key, n):
        fmt = AddressFormat.parse(fmt)
        stride = 32 if fmt == AddressFormat.P2tr else 20
        buf = C.create_string_buffer(stride * n)
        self._check(self._lib.vgen_cuda_digest_batch(self._ctx, int(fmt), _be(start_key), n, buf, stride))
        return buf.raw

    def set_pattern(self, pattern, fmt):
        self._check(self._lib.vgen_cuda_set_pattern(self._ctx, int(AddressFormat.parse(fmt)), pattern.original().encode(),
                                                    int(pattern.is_case_insensitive())))
        return self._lib.vgen_cuda_matcher_kind(self._ctx).decode()

    def matcher_kind(self):
        return self._lib.vgen_cuda_matcher_kind(self._ctx).decode()

    def lane_count(self):
        return int(self._lib.vgen_cuda_lane_count(self._ctx))

    def tap_bits(self):
        """Geometry of the P2TR fixed-base table built so far (0, a window width, or a signed-geometry word; see include/vgen_cuda.h)."""
        return int(self._lib.vgen_cuda_tap_bits(self._ctx))

    def launch_count(self):
        return int(self._lib.vgen_cuda_launch_count(self._ctx))

    def stats(self):
        """dict of the context's counters (vgen_cuda_stats): devices, failed_devices, requeued_windows, overflow_rescans, hit_cap, confirm_threads."""
        a = (C.c_uint64 * 8)()
        self._lib.vgen_cuda_stats(self._ctx, a)
        return dict(devices=a[0], failed_devices=a[1], requeued_windows=a[2], overflow_rescans=a[3], hit_cap=a[4], confirm_threads=a[5],
                    guarded_buffers=a[6], damaged_guards=a[7])

    def io_bytes(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._lib.vgen_cuda_io_bytes(self._ctx, C.byref(a), C.byref(b))
        return a.value, b.value

    def scan_raw(self, start, end, max_hits=0, stop=None, progress=None, hits_cap=4096, fetch_all=True):
        """vgen_cuda_scan: returns ([(key_int, digest_bytes, lane)], keys_scanned). All hits are returned: those beyond the
        first `hits_cap` are read back with vgen_cuda_scan_fetch_hits."""
        if getattr(self, "_hitbuf_cap", 0) < hits_cap:
            self._hitbuf = (_cabi.VgenHit * hits_cap)()
            self._hitbuf_cap = hits_cap
        hits = self._hitbuf
        n_hits = C.c_uint64(0)
        scanned = C.c_uint64(0)
        cb = _cabi.PROGRESS_FN(lambda k, u: progress(k)) if progress else C.cast(None, _cabi.PROGRESS_FN)
        stop_p = C.cast(stop, C.POINTER(C.c_int32)) if stop is not None else None
        rc = self._lib.vgen_cuda_scan(self._ctx, _be(start), None if end is None else _be(end), int(max_hits), stop_p, cb, None,
                                      hits, self._hitbuf_cap, C.byref(n_hits), C.byref(scanned))
        self._check(rc)

        def conv(buf, n):
            return [(int.from_bytes(bytes(h.key_be), "big"), bytes(h.digest)[:h.digest_len], h.device) for h in buf[:n]]
        out = conv(hits, n_hits.value)
        total = int(self._lib.vgen_cuda_scan_total_hits(self._ctx)) if fetch_all else 0
        while len(out) < total:
            n = int(self._lib.vgen_cuda_scan_fetch_hits(self._ctx, len(out), hits, self._hitbuf_cap))
            if n == 0:
                break
            out += conv(hits, n)
        return out, scanned.value

    def fetch_text(self, total):
        """"<key hex> <address> <wif>\\n" for every hit of the last scan (vgen_cuda_scan_fetch_text), copied out of the library into
        ONE caller-owned buffer. Returns (bytes-like, hits)."""
        if total <= 0:
            return b"", 0
        # fast path: the scan's confirm workers already built every record; one copy out of the library's buffer
        ptr, ln = C.c_void_p(), C.c_uint64(0)
        if self._lib.vgen_cuda_scan_text(self._ctx, C.cast(C.byref(ptr), C.POINTER(C.c_char_p)), C.byref(ln)):
            if ln.value >= _BIG_TEXT:
                # tens of MB: a new bytes object filled by the library on several threads instead of one memcpy here
                obj = _bytes_new(None, ln.value)
                nw = C.c_uint64(0)
                nb = int(self._lib.vgen_cuda_scan_fetch_text(self._ctx, 0, total, _bytes_addr(obj), ln.value + 1, C.byref(nw)))
                if nb == ln.value and nw.value == total:
                    return obj, total
            return C.string_at(ptr, ln.value), total
        cap = min(total * 227 + 1, 1 << 28)               # 227 = the library's bound for one record
        store = bytearray(cap)
        buf = (C.c_char * cap).from_buffer(store)
        n = C.c_uint64(0)
        got = pos = 0
        while got < total:
            room = C.cast(C.byref(buf, pos), C.c_void_p)
            nb = int(self._lib.vgen_cuda_scan_fetch_text(self._ctx, got, total - got, room, cap - pos, C.byref(n)))
            if nb <= 0 or n.value == 0:
                if pos == 0 or cap - pos >= 228:
                    break
                # buffer full (more than 2^28 bytes of records): continue in a fresh one
                more, g2 = self._fetch_text_rest(got, total)
                del buf
                return bytes(memoryview(store)[:pos]) + more, got + g2
            pos += nb
            got += n.value
        del buf
        return memoryview(store)[:pos], got

    def _fetch_text_rest(self, got, total):
        chunks = []
        cap = 1 << 26
        buf = C.create_string_buffer(cap)
        n = C.c_uint64(0)
        g0 = got
        while got < total:
            nb = int(self._lib.vgen_cuda_scan_fetch_text(self._ctx, got, total - got, buf, cap, C.byref(n)))
            if nb <= 0 or n.value == 0:
                break
            chunks.append(C.string_at(buf, nb))
            got += n.value
        return b"".join(chunks), got - g0

    def fetch_records(self, total):
        """(key_hex[], address[], wif[]) of every hit of the last scan, formatted by the library (vgen_cuda_scan_fetch_text)."""
        toks, got = [], 0
        cap = 1 << 22
        buf = C.create_string_buffer(cap)
        n = C.c_uint64(0)
        while got < total:
            nb = int(self._lib.vgen_cuda_scan_fetch_text(self._ctx, got, total - got, buf, cap, C.byref(n)))
            if nb <= 0 or n.value == 0:
                break
            toks += C.string_at(buf, nb).decode("ascii").split()
            got += n.value
        return toks[0::3], toks[1::3], toks[2::3]

    def bench_resident(self, start, n_windows):
        ms = C.c_float(0)
        keys = C.c_uint64(0)
        launches = C.c_uint64(0)
        self._check(self._lib.vgen_cuda_bench_resident(self._ctx, _be(start), n_windows, C.byref(ms), C.byref(keys), C.byref(launches)))
        return ms.value, keys.value, launches.value


def plan_windows(start, end, global_lanes, lane, sm_count=148, keys_per_launch_hint=0, can_end_early=False):
    """[(first_key, n_keys)] that global lane `lane` of `global_lanes` scans in [start, end] (vgen_plan_windows; no GPU needed)."""
    L = _cabi.lib()
    n = L.vgen_plan_windows(_be(start), _be(end), sm_count, keys_per_launch_hint, global_lanes, lane, int(can_end_early), None, None, 0)
    if n < 0:
        raise _cabi.VgenError(int(n), "vgen_plan_windows")
    keys = C.create_string_buffer(32 * max(n, 1))
    lens = (C.c_uint64 * max(n, 1))()
    L.vgen_plan_windows(_be(start), _be(end), sm_count, keys_per_launch_hint, global_lanes, lane, int(can_end_early), keys, lens, n)
    raw = keys.raw
    return [(int.from_bytes(raw[32 * i:32 * i + 32], "big"), int(lens[i])) for i in range(n)]


# bytes objects the library fills in place (PyBytes_FromStringAndSize(NULL, n): n bytes, uninitialised, plus the trailing NUL)
_BIG_TEXT = 1 << 20
_bytes_new = C.pythonapi.PyBytes_FromStringAndSize
_bytes_new.restype, _bytes_new.argtypes = C.py_object, [C.c_void_p, C.c_ssize_t]
_bytes_addr = C.pythonapi.PyBytes_AsString
_bytes_addr.restype, _bytes_addr.argtypes = C.c_void_p, [C.py_object]


class DigestBatch(collections.abc.Sequence):
    """Digests of consecutive keys as ONE bytes object (`raw`, `stride` bytes per key) with sequence access to the per-key
    digests: what the reference returns as Vec<[u8; 20]> (src/gpu.rs:644-650) without building one Python object per key."""

    def __init__(self, raw, stride):
        self.raw, self.stride = raw, int(stride)

    def __len__(self):
        return len(self.raw) // self.stride

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError(i)
        return self.raw[self.stride * i:self.stride * (i + 1)]

    def __bytes__(self):
        return self.raw

    def __eq__(self, other):
        if isinstance(other, DigestBatch):
            return self.raw == other.raw and self.stride == other.stride
        return list(self) == list(other)


class StopFlag:
    """Arc<AtomicBool> of the reference (src/gpu.rs:924): a 32-bit flag the library polls after every window."""

    def __init__(self):
        self._v = C.c_int32(0)

    def set(self):
        self._v.value = 1

    def is_set(self):
        return self._v.value != 0

    def _ptr(self):
        return C.pointer(self._v)


def scan_gpu_with_runner(pattern, config, progress_cb, stop, runner):
    """src/gpu.rs:920-1125: sequential scan from config.start (or a random valid key) to config.end, first `count`
    matches in ascending key order. operations = exact valid keys tested (Appendix A of SURVEY.md).
    `count` has the library's meaning: the scan ends once `count` matches are found, so count == 0 returns at once with no
    matches and no operations (src/gpu.rs:981-996: `found >= config.count` holds before the first dispatch); "no limit" is
    UNLIMITED = usize::MAX, which is what the reference's CLI substitutes for `--count 0` (src/lib.rs:524)."""
    if not isinstance(pattern, Pattern):
        pattern = Pattern(pattern)
    fmt = AddressFormat.parse(config.format)
    if config.start is not None:
        start = int(config.start)
        if start == 0 and config.end is None:
            raise ValueError("Invalid start key")                         # src/gpu.rs:903
    else:
        start = 1 + secrets.randbelow(SECP256K1_N - 1)                    # src/gpu.rs:933-945
    end = None if config.end is None else int(config.end)
    runner.set_pattern(pattern, fmt)
    t0 = time.perf_counter()
    count = int(config.count)
    if count <= 0:
        return ScanResult(matches=MatchList(b"", 0, fmt), operations=0, elapsed_secs=time.perf_counter() - t0)
    max_hits = 0 if count >= UNLIMITED else count                         # C ABI: max_hits 0 = unlimited
    # first call with a 1-record buffer: the confirmed hits stay in the context and are formatted in bulk below
    _, scanned = runner.scan_raw(start, end, max_hits=max_hits, stop=stop._ptr() if stop is not None else None, progress=progress_cb,
                                 hits_cap=1, fetch_all=False)
    total = int(runner._lib.vgen_cuda_scan_total_hits(runner._ctx))
    text, got = runner.fetch_text(total)
    matches = MatchList(text, got, fmt)
    elapsed = time.perf_counter() - t0
    return ScanResult(matches=matches, operations=scanned, elapsed_secs=elapsed)


def scan_gpu_p2tr_with_runner(pattern, config, progress_cb, stop, runner):
    """src/gpu.rs:1168-1343. Here P2TR runs through the same entry point as every other format (both stages on the GPU)."""
    import dataclasses
    return scan_gpu_with_runner(pattern, dataclasses.replace(config, format=AddressFormat.P2tr), progress_cb, stop, runner)


def scan_gpu_p2tr(pattern, config, progress_cb=None, stop=None):
    """src/gpu.rs:1346-1366."""
    import dataclasses
    return scan_gpu(pattern, dataclasses.replace(config, format=AddressFormat.P2tr), progress_cb, stop)


def scan_gpu(pattern, config, progress_cb=None, stop=None):
    """src/gpu.rs:1127-1149: one-shot wrapper that builds its own runner."""
    runner = GpuRunner(config.gpu_batch_size or 0, GpuBackend.Auto)
    try:
        return scan_gpu_with_runner(pattern, config, progress_cb, stop, runner)
    finally:
        runner.close()
