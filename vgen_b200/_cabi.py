"""ctypes binding of libvgen_cuda.so (include/vgen_cuda.h). Fails loudly when the library is missing:
there is no CPU fallback on the product path (north_star; reference fallback branches
src/lib.rs:727-746,1185-1197 become hard errors)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# VGEN_CUDA_LIB: development override to load a differently-built library (kernel tuning variants)
LIB_PATH = os.environ.get("VGEN_CUDA_LIB") or os.path.join(_HERE, "libvgen_cuda.so")

VGEN_OK = 0
ERR_NAMES = {-1: "invalid argument", -2: "CUDA error", -3: "invalid pattern", -4: "invalid range", -5: "buffer too small"}


class VgenHit(C.Structure):
    _fields_ = [("key_be", C.c_uint8 * 32), ("digest", C.c_uint8 * 32), ("digest_len", C.c_uint32), ("device", C.c_uint32)]


PROGRESS_FN = C.CFUNCTYPE(None, C.c_uint64, C.c_void_p)

# every symbol include/vgen_cuda.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "vgen_cuda_device_count": (C.c_int, []),
    "vgen_cuda_device_name": (C.c_int, [C.c_int, C.c_char_p, C.c_size_t]),
    "vgen_cuda_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "vgen_cuda_destroy": (None, [C.c_void_p]),
    "vgen_cuda_last_error": (C.c_char_p, [C.c_void_p]),
    "vgen_cuda_set_pattern": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_int]),
    "vgen_cuda_matcher_kind": (C.c_char_p, [C.c_void_p]),
    "vgen_cuda_set_shard": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32]),
    "vgen_cuda_set_stream": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "vgen_cuda_scan": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_int32), PROGRESS_FN, C.c_void_p,
                                 C.POINTER(VgenHit), C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "vgen_cuda_scan_total_hits": (C.c_uint64, [C.c_void_p]),
    "vgen_cuda_scan_fetch_hits": (C.c_uint64, [C.c_void_p, C.c_uint64, C.POINTER(VgenHit), C.c_uint64]),
    "vgen_cuda_scan_fetch_records": (C.c_uint64, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "vgen_cuda_scan_fetch_text": (C.c_int64, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]),
    "vgen_cuda_scan_text": (C.c_int, [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_uint64)]),
    "vgen_cuda_hash160_batch": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p]),
    "vgen_cuda_digest_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_uint64, C.c_void_p, C.c_uint32]),
    # allocation-free hash160 batch (static device work area; out is a DEVICE pointer): device, start, n, out20_dev, stream, err, err_cap
    "vgen_cuda_hash160_batch_static": (C.c_int, [C.c_int, C.c_char_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]),
    "vgen_cuda_bench_resident": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "vgen_cuda_launch_count": (C.c_uint64, [C.c_void_p]),
    "vgen_cuda_window_keys": (C.c_uint64, [C.c_void_p]),
    "vgen_cuda_lane_count": (C.c_int, [C.c_void_p]),
    "vgen_cuda_tap_bits": (C.c_int, [C.c_void_p]),
    "vgen_cuda_stats": (None, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "vgen_cuda_io_bytes": (None, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "vgen_format_address": (C.c_int, [C.c_int, C.c_char_p, C.c_char_p, C.c_size_t]),
    "vgen_format_wif": (C.c_int, [C.c_int, C.c_char_p, C.c_char_p, C.c_size_t]),
    "vgen_pattern_check": (C.c_int, [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]),
    "vgen_pattern_is_match": (C.c_int, [C.c_char_p, C.c_int, C.c_char_p]),
    "vgen_plan_windows": (C.c_int64, [C.c_char_p, C.c_char_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]),
    "vgen_pattern_dfa": (C.c_int, [C.c_int, C.c_char_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_uint32)]),
    "vgen_pattern_lowering": (C.c_int, [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]),
    "vgen_pattern_validate_charset": (C.c_int, [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]),
    "vgen_pattern_estimate_difficulty": (C.c_uint64, [C.c_int, C.c_char_p, C.c_int]),
    "vgen_provider_resolve": (C.c_int, [C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int), C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_char_p, C.c_size_t]),
    "vgen_provider_pattern": (C.c_int, [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]),
    "vgen_parse_key": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
}

_lib = None


class VgenError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (ERR_NAMES.get(code, "error"), code, msg))
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libvgen_cuda.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` or "
                              "`make -C vgen_b200/csrc`); vgen_b200 has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)          # AttributeError if the library lacks a declared symbol
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib
