"""vgen_b200 — B200 (sm_100a) keygen-and-match engine behind vgen's GPU scan surface.

Python host mirror of the reference's crate-root re-exports (src/lib.rs:9-12) over the C ABI of libvgen_cuda.so
(include/vgen_cuda.h). There is no CPU fallback: importing works without a GPU (for host-side logic and tests), but
every scan entry point needs the built library and a B200."""
from . import _cabi
from .address import AddressFormat, GeneratedAddress, address_from_digest, wif_from_key
from .gpu import DEFAULT_BATCH_SIZE, UNLIMITED, DigestBatch, GpuBackend, GpuRunner, StopFlag, plan_windows, scan_gpu, scan_gpu_p2tr, scan_gpu_p2tr_with_runner, scan_gpu_with_runner
from .pattern import Pattern
from .scanner import ScanConfig, ScanResult

__all__ = ["AddressFormat", "GeneratedAddress", "GpuBackend", "GpuRunner", "StopFlag", "Pattern", "ScanConfig", "ScanResult",
           "scan_gpu", "scan_gpu_with_runner", "scan_gpu_p2tr", "scan_gpu_p2tr_with_runner", "plan_windows", "DEFAULT_BATCH_SIZE", "UNLIMITED", "DigestBatch", "address_from_digest", "wif_from_key"]
__version__ = "0.1.0"
