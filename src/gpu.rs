//! GPU scan path of vgen on NVIDIA B200 (sm_100a): a thin FFI binding to `libvgen_cuda.so` (C ABI: `include/vgen_cuda.h`).
//!
//! Drop-in replacement for the reference's `src/gpu.rs` (wgpu + WGSL, 1,460 lines): every `pub` item the crate and its
//! callers use keeps its name and signature —
//!   `GpuBackend` (+ `fallback_order`, `name`, `Display`)            reference src/gpu.rs:18-63
//!   `DEFAULT_BATCH_SIZE`                                            :83
//!   `GpuRunner { pub batch_size, pub device_name }`, `backend()`     :116-136
//!   `GpuRunner::new`                                                :138
//!   `dispatch` / `await_result` / `generate_batch`                  :535 / :602 / :660
//!   `dispatch_p2tr` / `await_result_p2tr` / `dispatch_batch_hash`   :667 / :750 / :809
//!   `scan_gpu_with_runner` / `scan_gpu`                             :920 / :1127
//!   `scan_gpu_p2tr_with_runner` / `scan_gpu_p2tr`                   :1168 / :1346
//! — so `src/lib.rs` (`:717,835-853,1163-1184`) and `benches/gpu_bench.rs:33-48` compile against it unchanged, except for
//! the two items that exposed wgpu types (`GpuBackend::to_wgpu_backends`, `is_software_adapter(&wgpu::AdapterInfo)`; see
//! INTEGRATION.md for the two call sites).
//!
//! What is different underneath: address encoding and regex matching run on the GPU, only confirmed hits cross PCIe;
//! Ethereum and P2TR scan on the GPU too; there is NO CPU fallback (a missing library or device is an `Err`).
//!
//! This file cannot be compiled in the build image (no cargo/rustc there). The identical call sequence is exercised by a C
//! client of the same header, `tests/cabi_harness.c`, and by the Python mirror `vgen_b200/gpu.py`.
#![allow(unsafe_code)] // Cargo.toml: `unsafe_code = "forbid"` becomes "deny"; FFI needs it in this module only

use std::ffi::{c_char, c_int, c_void, CStr, CString};
use std::sync::atomic::{AtomicBool, AtomicI32, Ordering};
use std::sync::{Arc, Mutex};
use std::thread::JoinHandle;
use std::time::Instant;

use anyhow::{anyhow, Result};
use clap::ValueEnum;
use num_bigint::BigUint;
use rand::rngs::StdRng;
use rand::{Rng, SeedableRng};

use crate::address::{AddressFormat, GeneratedAddress};
use crate::{Pattern, ProgressCallback, ScanConfig, ScanResult};

// ---------------------------------------------------------------------------------------------------------------------
// C ABI (include/vgen_cuda.h)
// ---------------------------------------------------------------------------------------------------------------------

#[repr(C)]
pub struct VgenCtx {
    _opaque: [u8; 0],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct VgenHit {
    pub key_be: [u8; 32],
    pub digest: [u8; 32],
    pub digest_len: u32,
    pub device: u32,
}

type ProgressFn = Option<extern "C" fn(u64, *mut c_void)>;

/// `format` argument of `vgen_cuda_digest_batch` only: the untweaked even-Y internal key (what `await_result_p2tr` returns).
const VGEN_P2TR_INTERNAL_KEY: c_int = 6;

#[link(name = "vgen_cuda")]
extern "C" {
    fn vgen_cuda_device_count() -> c_int;
    fn vgen_cuda_device_name(device: c_int, buf: *mut c_char, len: usize) -> c_int;
    fn vgen_cuda_create(devices: *const c_int, n_devices: c_int, keys_per_launch_hint: u32, out: *mut *mut VgenCtx) -> c_int;
    fn vgen_cuda_destroy(ctx: *mut VgenCtx);
    fn vgen_cuda_last_error(ctx: *const VgenCtx) -> *const c_char;
    fn vgen_cuda_set_pattern(ctx: *mut VgenCtx, format: c_int, pattern: *const c_char, case_insensitive: c_int) -> c_int;
    fn vgen_cuda_scan(
        ctx: *mut VgenCtx,
        start_be: *const u8,
        end_be: *const u8,
        max_hits: u64,
        stop: *const i32,
        progress: ProgressFn,
        user: *mut c_void,
        hits_out: *mut VgenHit,
        hits_cap: u64,
        n_hits: *mut u64,
        keys_scanned: *mut u64,
    ) -> c_int;
    fn vgen_cuda_scan_total_hits(ctx: *const VgenCtx) -> u64;
    fn vgen_cuda_scan_fetch_hits(ctx: *const VgenCtx, offset: u64, out: *mut VgenHit, cap: u64) -> u64;
    fn vgen_cuda_hash160_batch(ctx: *mut VgenCtx, start_be: *const u8, n: u64, out20: *mut u8) -> c_int;
    /// Allocation-free variant (static device work area, at most 2^20 keys per call, `out20_dev` is DEVICE memory). Not used by
    /// `GpuRunner`: it exists for callers under CUDA graph capture. See include/vgen_cuda.h.
    #[allow(dead_code)]
    fn vgen_cuda_hash160_batch_static(device: c_int, start_be: *const u8, n: u32, out20_dev: *mut u8, cuda_stream: *mut c_void,
                                      err: *mut c_char, err_cap: usize) -> c_int;
    fn vgen_cuda_digest_batch(ctx: *mut VgenCtx, format: c_int, start_be: *const u8, n: u64, out: *mut u8, stride: u32) -> c_int;
    fn vgen_cuda_window_keys(ctx: *const VgenCtx) -> u64;
    fn vgen_format_address(format: c_int, digest: *const u8, out: *mut c_char, cap: usize) -> c_int;
    fn vgen_format_wif(format: c_int, key_be: *const u8, out: *mut c_char, cap: usize) -> c_int;
}

fn last_error(ctx: *const VgenCtx) -> String {
    // SAFETY: the library returns a NUL-terminated string that stays valid until the next call on this context / thread
    unsafe { CStr::from_ptr(vgen_cuda_last_error(ctx)) }.to_string_lossy().into_owned()
}

fn format_id(format: AddressFormat) -> c_int {
    match format {
        AddressFormat::P2pkh => 0,
        AddressFormat::P2pkhUncompressed => 1,
        AddressFormat::P2wpkh => 2,
        AddressFormat::P2shP2wpkh => 3,
        AddressFormat::P2tr => 4,
        AddressFormat::Ethereum => 5,
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// GpuBackend (reference src/gpu.rs:18-63)
// ---------------------------------------------------------------------------------------------------------------------

/// The wgpu variants stay so that `--backend vulkan|metal|dx12|gl` still parses; asking for one of them is a clear error
/// in `GpuRunner::new` (this build drives B200 GPUs through CUDA only).
#[derive(Debug, Clone, Copy, PartialEq, Eq, ValueEnum, Default)]
pub enum GpuBackend {
    #[default]
    Auto,
    Cuda,
    Vulkan,
    Metal,
    Dx12,
    Gl,
}

impl GpuBackend {
    /// Backends tried under `Auto`, in order. CUDA is the only one.
    pub fn fallback_order() -> &'static [GpuBackend] {
        &[GpuBackend::Cuda]
    }

    pub fn name(self) -> &'static str {
        match self {
            GpuBackend::Auto => "auto",
            GpuBackend::Cuda => "CUDA",
            GpuBackend::Vulkan => "Vulkan",
            GpuBackend::Metal => "Metal",
            GpuBackend::Dx12 => "DX12",
            GpuBackend::Gl => "OpenGL",
        }
    }
}

impl std::fmt::Display for GpuBackend {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        write!(f, "{}", self.name())
    }
}

/// `list-gpus` (src/lib.rs:979-1036) used this to hide llvmpipe & co. A CUDA device is never a software adapter.
pub(crate) fn is_software_adapter(_device_name: &str) -> bool {
    false
}

/// Names of the CUDA devices the library sees (replaces the wgpu adapter enumeration of `list-gpus`).
pub fn list_devices() -> Vec<String> {
    // SAFETY: plain value call
    let n = unsafe { vgen_cuda_device_count() };
    (0..n)
        .map(|i| {
            let mut buf = [0 as c_char; 256];
            // SAFETY: buf is writable for its whole length
            unsafe { vgen_cuda_device_name(i, buf.as_mut_ptr(), buf.len()) };
            // SAFETY: the library NUL-terminates
            unsafe { CStr::from_ptr(buf.as_ptr()) }.to_string_lossy().into_owned()
        })
        .collect()
}

/// Keys per `dispatch` frame / `generate_batch` call (reference :83). For scans it is only a keys-per-launch hint:
/// `GpuRunner::new(0, ..)` lets the library fill the GPU (77.6 M keys per launch on a B200).
pub const DEFAULT_BATCH_SIZE: u32 = 512 * 1024;

// ---------------------------------------------------------------------------------------------------------------------
// GpuRunner (reference src/gpu.rs:116-888)
// ---------------------------------------------------------------------------------------------------------------------

enum FrameOut {
    Hash160(Vec<[u8; 20]>),
    XOnly(Vec<[u8; 32]>),
}

#[derive(Default)]
struct Frame {
    job: Mutex<Option<JoinHandle<Result<FrameOut>>>>,
    batch_start_key: Mutex<[u8; 32]>,
}

struct Ctx(*mut VgenCtx);
// SAFETY: the context is only ever used under `GpuRunner::ctx` 's mutex (the library allows one job at a time per context)
unsafe impl Send for Ctx {}

pub struct GpuRunner {
    ctx: Arc<Mutex<Ctx>>,
    frames: Vec<Frame>, // two frames in flight, as the reference (:399)
    pub batch_size: u32,
    pub device_name: String,
    backend: GpuBackend,
    /// keys one scan launch covers on this device (informational)
    pub window_keys: u64,
}

impl Drop for GpuRunner {
    fn drop(&mut self) {
        for f in &self.frames {
            if let Some(h) = f.job.lock().unwrap().take() {
                let _ = h.join();
            }
        }
        let ctx = self.ctx.lock().unwrap();
        // SAFETY: created by vgen_cuda_create, destroyed exactly once
        unsafe { vgen_cuda_destroy(ctx.0) };
    }
}

impl GpuRunner {
    pub fn backend(&self) -> GpuBackend {
        self.backend
    }

    /// reference :138. `batch_size` = keys per dispatch frame and keys-per-launch hint of scans (0: fill the GPU).
    /// Errors instead of falling back: unknown backend, no library, no sm_100 device.
    pub async fn new(batch_size: u32, backend: GpuBackend) -> Result<Self> {
        Self::with_devices(batch_size, backend, &[0])
    }

    /// One context over several GPUs (`--devices 0,1,..`): one host thread per device inside the library, hits merged.
    pub fn with_devices(batch_size: u32, backend: GpuBackend, devices: &[i32]) -> Result<Self> {
        if !matches!(backend, GpuBackend::Auto | GpuBackend::Cuda) {
            return Err(anyhow!("backend {backend} is not available: this build drives B200 GPUs through CUDA only"));
        }
        let devs: Vec<c_int> = devices.iter().map(|&d| d as c_int).collect();
        let mut ctx: *mut VgenCtx = std::ptr::null_mut();
        // SAFETY: devs outlives the call, ctx is a valid out-pointer
        let rc = unsafe { vgen_cuda_create(devs.as_ptr(), devs.len() as c_int, batch_size, &mut ctx) };
        if rc != 0 {
            return Err(anyhow!("GPU initialization failed: {}", last_error(std::ptr::null())));
        }
        let mut name = [0 as c_char; 256];
        // SAFETY: name is writable for its whole length
        unsafe { vgen_cuda_device_name(devs.first().copied().unwrap_or(0), name.as_mut_ptr(), name.len()) };
        // SAFETY: NUL-terminated by the library
        let device_name = unsafe { CStr::from_ptr(name.as_ptr()) }.to_string_lossy().into_owned();
        // SAFETY: ctx is valid
        let window_keys = unsafe { vgen_cuda_window_keys(ctx) };
        Ok(Self {
            ctx: Arc::new(Mutex::new(Ctx(ctx))),
            frames: vec![Frame::default(), Frame::default()],
            batch_size: if batch_size == 0 { DEFAULT_BATCH_SIZE } else { batch_size },
            device_name,
            backend: GpuBackend::Cuda,
            window_keys,
        })
    }

    fn spawn_frame(&self, start_key: [u8; 32], frame_index: usize, p2tr: bool) -> Result<()> {
        let frame = self.frames.get(frame_index).ok_or_else(|| anyhow!("frame index {frame_index} out of range"))?;
        let mut guard = frame.job.lock().unwrap();
        if let Some(prev) = guard.take() {
            let _ = prev.join(); // a frame that was never awaited is dropped, as the reference unmaps its staging buffer
        }
        *frame.batch_start_key.lock().unwrap() = start_key;
        let ctx = Arc::clone(&self.ctx);
        let n = self.batch_size as usize;
        *guard = Some(std::thread::spawn(move || -> Result<FrameOut> {
            let ctx = ctx.lock().unwrap();
            if p2tr {
                let mut out = vec![[0u8; 32]; n];
                // SAFETY: out holds n * 32 writable bytes
                let rc = unsafe { vgen_cuda_digest_batch(ctx.0, VGEN_P2TR_INTERNAL_KEY, start_key.as_ptr(), n as u64, out.as_mut_ptr().cast(), 32) };
                if rc != 0 {
                    return Err(anyhow!("{}", last_error(ctx.0)));
                }
                Ok(FrameOut::XOnly(out))
            } else {
                let mut out = vec![[0u8; 20]; n];
                // SAFETY: out holds n * 20 writable bytes
                let rc = unsafe { vgen_cuda_hash160_batch(ctx.0, start_key.as_ptr(), n as u64, out.as_mut_ptr().cast()) };
                if rc != 0 {
                    return Err(anyhow!("{}", last_error(ctx.0)));
                }
                Ok(FrameOut::Hash160(out))
            }
        }));
        Ok(())
    }

    fn join_frame(&self, frame_index: usize) -> Result<(FrameOut, [u8; 32])> {
        let frame = self.frames.get(frame_index).ok_or_else(|| anyhow!("frame index {frame_index} out of range"))?;
        let job = frame.job.lock().unwrap().take().ok_or_else(|| anyhow!("frame {frame_index} was not dispatched"))?;
        let out = job.join().map_err(|_| anyhow!("GPU frame worker panicked"))??;
        let start_key = *frame.batch_start_key.lock().unwrap();
        Ok((out, start_key))
    }

    /// reference :535 — hash160 of `batch_size` sequential keys from `start_key`, computed in the background.
    pub fn dispatch(&self, start_key: [u8; 32], frame_index: usize) -> Result<()> {
        self.spawn_frame(start_key, frame_index, false)
    }

    /// reference :602
    pub async fn await_result(&self, frame_index: usize) -> Result<(Vec<[u8; 20]>, [u8; 32])> {
        match self.join_frame(frame_index)? {
            (FrameOut::Hash160(v), k) => Ok((v, k)),
            _ => Err(anyhow!("frame {frame_index} holds a P2TR result")),
        }
    }

    /// reference :660
    pub async fn generate_batch(&self, start_key: [u8; 32]) -> Result<Vec<[u8; 20]>> {
        self.dispatch(start_key, 0)?;
        let (res, _) = self.await_result(0).await?;
        Ok(res)
    }

    /// reference :667 — x-only even-Y INTERNAL keys (before the TapTweak), 32 bytes each, big-endian.
    pub fn dispatch_p2tr(&self, start_key: [u8; 32], frame_index: usize) -> Result<()> {
        self.spawn_frame(start_key, frame_index, true)
    }

    /// reference :750
    pub async fn await_result_p2tr(&self, frame_index: usize) -> Result<(Vec<[u8; 32]>, [u8; 32])> {
        match self.join_frame(frame_index)? {
            (FrameOut::XOnly(v), k) => Ok((v, k)),
            _ => Err(anyhow!("frame {frame_index} holds a hash160 result")),
        }
    }

    /// reference :809 — same result as `dispatch` (the library always uses batch inversion).
    pub fn dispatch_batch_hash(&self, start_key: [u8; 32], frame_index: usize) -> Result<()> {
        self.dispatch(start_key, frame_index)
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Scans (reference src/gpu.rs:920-1366)
// ---------------------------------------------------------------------------------------------------------------------

fn biguint_to_bytes32(n: &BigUint) -> [u8; 32] {
    let bytes = n.to_bytes_be();
    let mut result = [0u8; 32];
    let start = 32usize.saturating_sub(bytes.len());
    result[start..].copy_from_slice(&bytes[bytes.len().saturating_sub(32)..]);
    result
}

/// secp256k1 group order
const ORDER: [u8; 32] = [
    0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFE, 0xBA, 0xAE, 0xDC, 0xE6, 0xAF, 0x48, 0xA0, 0x3B,
    0xBF, 0xD2, 0x5E, 0x8C, 0xD0, 0x36, 0x41, 0x41,
];

fn random_valid_key() -> [u8; 32] {
    // rejection sampling, as the reference (:933-945)
    let mut rng = StdRng::from_entropy();
    loop {
        let k: [u8; 32] = rng.gen();
        if k != [0u8; 32] && k < ORDER {
            break k;
        }
    }
}

struct ScanHooks<'a> {
    progress: Option<&'a ProgressCallback>,
    stop: Option<&'a AtomicBool>,
    flag: AtomicI32, // what the library polls after every window
}

extern "C" fn progress_trampoline(keys_scanned: u64, user: *mut c_void) {
    // SAFETY: `user` is the &ScanHooks that lives on scan_with_format's stack for the whole vgen_cuda_scan call; the library
    // calls this on the calling thread only
    let hooks = unsafe { &*(user as *const ScanHooks) };
    if let Some(cb) = hooks.progress {
        cb(keys_scanned);
    }
    if let Some(stop) = hooks.stop {
        if stop.load(Ordering::Relaxed) {
            hooks.flag.store(1, Ordering::Relaxed);
        }
    }
}

fn c_string(buf: &[c_char]) -> String {
    // SAFETY: the library NUL-terminates within the buffer
    unsafe { CStr::from_ptr(buf.as_ptr()) }.to_string_lossy().into_owned()
}

fn scan_with_format(
    pattern: &Pattern,
    config: &ScanConfig,
    format: AddressFormat,
    progress_cb: Option<ProgressCallback>,
    stop: Option<Arc<AtomicBool>>,
    runner: &GpuRunner,
) -> Result<ScanResult> {
    let t0 = Instant::now();
    // reference :981-996: `found >= config.count` is checked before the first dispatch
    if config.count == 0 || stop.as_ref().is_some_and(|s| s.load(Ordering::Relaxed)) {
        return Ok(ScanResult { matches: Vec::new(), operations: 0, elapsed_secs: t0.elapsed().as_secs_f64() });
    }
    let start = match &config.start {
        Some(s) => biguint_to_bytes32(s),
        None => random_valid_key(),
    };
    if config.start.is_some() && start == [0u8; 32] && config.end.is_none() {
        return Err(anyhow!("Invalid start key")); // reference :903 (SecretKey::from_slice rejects 0)
    }
    let end = config.end.as_ref().map(biguint_to_bytes32);
    let fmt = format_id(format);
    let pat = CString::new(pattern.original())?;
    let ctx = runner.ctx.lock().unwrap();

    // SAFETY: ctx and pat are valid for the call
    let rc = unsafe { vgen_cuda_set_pattern(ctx.0, fmt, pat.as_ptr(), c_int::from(pattern.is_case_insensitive())) };
    if rc != 0 {
        return Err(anyhow!("{}", last_error(ctx.0)));
    }

    let hooks = ScanHooks { progress: progress_cb.as_ref(), stop: stop.as_deref(), flag: AtomicI32::new(0) };
    let max_hits = if config.count == usize::MAX { 0 } else { config.count as u64 }; // C ABI: 0 = unlimited
    let mut first = [VgenHit { key_be: [0; 32], digest: [0; 32], digest_len: 0, device: 0 }; 64];
    let (mut n_hits, mut keys_scanned) = (0u64, 0u64);
    // SAFETY: every pointer is valid for the duration of the call; `hooks` outlives it; the stop flag is an i32 the library
    // only reads
    let rc = unsafe {
        vgen_cuda_scan(
            ctx.0,
            start.as_ptr(),
            end.as_ref().map_or(std::ptr::null(), |e| e.as_ptr()),
            max_hits,
            hooks.flag.as_ptr(),
            Some(progress_trampoline),
            &hooks as *const ScanHooks as *mut c_void,
            first.as_mut_ptr(),
            first.len() as u64,
            &mut n_hits,
            &mut keys_scanned,
        )
    };
    if rc != 0 {
        return Err(anyhow!("{}", last_error(ctx.0)));
    }
    let mut hits: Vec<VgenHit> = first[..n_hits as usize].to_vec();
    // SAFETY: ctx is valid
    let total = unsafe { vgen_cuda_scan_total_hits(ctx.0) };
    while (hits.len() as u64) < total {
        let mut more = vec![VgenHit { key_be: [0; 32], digest: [0; 32], digest_len: 0, device: 0 }; 4096];
        // SAFETY: `more` holds 4096 writable records
        let got = unsafe { vgen_cuda_scan_fetch_hits(ctx.0, hits.len() as u64, more.as_mut_ptr(), more.len() as u64) };
        if got == 0 {
            break;
        }
        more.truncate(got as usize);
        hits.extend(more);
    }
    drop(ctx);

    // address / WIF strings per match (the reference builds them for every key, :1030-1093; here only for hits)
    let matches = hits
        .iter()
        .map(|h| {
            let mut addr = [0 as c_char; 128];
            let mut wif = [0 as c_char; 128];
            // SAFETY: buffers are writable for their whole length; digest / key pointers are valid
            unsafe {
                vgen_format_address(fmt, h.digest.as_ptr(), addr.as_mut_ptr(), addr.len());
                vgen_format_wif(fmt, h.key_be.as_ptr(), wif.as_mut_ptr(), wif.len());
            }
            GeneratedAddress { address: c_string(&addr), wif: c_string(&wif), hex: hex::encode(h.key_be), format }
        })
        .collect();
    Ok(ScanResult { matches, operations: keys_scanned, elapsed_secs: t0.elapsed().as_secs_f64() })
}

/// reference :920 — same signature. Sequential scan from `config.start` (or a random valid key) to `config.end`; the first
/// `config.count` matches in ascending key order; `operations` = valid keys tested.
pub async fn scan_gpu_with_runner(
    pattern: &Pattern,
    config: &ScanConfig,
    progress_cb: Option<ProgressCallback>,
    stop: Option<Arc<AtomicBool>>,
    runner: Arc<GpuRunner>,
) -> Result<ScanResult> {
    scan_with_format(pattern, config, config.format, progress_cb, stop, &runner)
}

/// reference :1127
pub fn scan_gpu(
    pattern: &Pattern,
    config: &ScanConfig,
    progress_cb: Option<ProgressCallback>,
    stop: Option<Arc<AtomicBool>>,
) -> Result<ScanResult> {
    let rt = tokio::runtime::Builder::new_current_thread().enable_all().build()?;
    let batch_size = config.gpu_batch_size.unwrap_or(0); // 0: let the library fill the GPU
    let runner = Arc::new(rt.block_on(GpuRunner::new(batch_size, GpuBackend::Auto))?);
    rt.block_on(scan_gpu_with_runner(pattern, config, progress_cb, stop, runner))
}

/// reference :1168 — P2TR. Both stages (even-Y internal key, TapTweak + t*G) run on the device; the per-key host tweak of the
/// reference (:1288-1290) is gone.
pub async fn scan_gpu_p2tr_with_runner(
    pattern: &Pattern,
    config: &ScanConfig,
    progress_cb: Option<ProgressCallback>,
    stop: Option<Arc<AtomicBool>>,
    runner: Arc<GpuRunner>,
) -> Result<ScanResult> {
    scan_with_format(pattern, config, AddressFormat::P2tr, progress_cb, stop, &runner)
}

/// reference :1346
pub fn scan_gpu_p2tr(
    pattern: &Pattern,
    config: &ScanConfig,
    progress_cb: Option<ProgressCallback>,
    stop: Option<Arc<AtomicBool>>,
) -> Result<ScanResult> {
    let rt = tokio::runtime::Builder::new_current_thread().enable_all().build()?;
    let batch_size = config.gpu_batch_size.unwrap_or(0);
    let runner = Arc::new(rt.block_on(GpuRunner::new(batch_size, GpuBackend::Auto))?);
    rt.block_on(scan_gpu_p2tr_with_runner(pattern, config, progress_cb, stop, runner))
}

#[cfg(test)]
mod tests {
    use super::*;

    #[test]
    fn test_gpu_backend_fallback_order() {
        assert_eq!(GpuBackend::fallback_order(), &[GpuBackend::Cuda]);
    }

    #[test]
    fn test_gpu_backend_display() {
        assert_eq!(GpuBackend::Auto.to_string(), "auto");
        assert_eq!(GpuBackend::Cuda.to_string(), "CUDA");
        assert_eq!(GpuBackend::Vulkan.to_string(), "Vulkan");
        assert_eq!(GpuBackend::Metal.to_string(), "Metal");
        assert_eq!(GpuBackend::Dx12.to_string(), "DX12");
        assert_eq!(GpuBackend::Gl.to_string(), "OpenGL");
    }

    #[test]
    fn test_is_software_adapter() {
        assert!(!is_software_adapter("NVIDIA B200"));
    }

    #[test]
    fn test_biguint_to_bytes32() {
        let mut want = [0u8; 32];
        want[23] = 1; // 2^64
        assert_eq!(biguint_to_bytes32(&(BigUint::from(1u8) << 64)), want);
        assert_eq!(biguint_to_bytes32(&BigUint::from(0u8)), [0u8; 32]);
    }

    #[test]
    fn test_unavailable_backend_is_an_error() {
        let rt = tokio::runtime::Builder::new_current_thread().build().unwrap();
        assert!(rt.block_on(GpuRunner::new(0, GpuBackend::Vulkan)).is_err());
    }
}
