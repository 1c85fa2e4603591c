// cuda_index.rs — the device handle the three EncodedVectors types gain with the `cuda` feature, and the stop-flag
// bridge.  SYNTAX-REVIEWED ONLY (no Rust toolchain in the build image).
use crate::cuda_ffi::*;
use crate::EncodingError;
use std::sync::atomic::{AtomicI32, Ordering};
use std::sync::Arc;

/// One context per host thread that should run concurrently: calls on one context are serialised inside the
/// library (one stream, one scratch set).  Created with stream = NULL, so the context owns its stream and
/// consecutive *_dev scans are chained automatically (include/qz_cuda.h, qz_ctx_set_scan_chaining).
pub struct CudaContext(pub *mut qz_ctx);
unsafe impl Send for CudaContext {}

impl CudaContext {
    pub fn new(device: i32) -> Result<Self, EncodingError> {
        let mut ctx = std::ptr::null_mut();
        let st = unsafe { qz_ctx_create(device, std::ptr::null_mut(), &mut ctx) };
        if st != 0 {
            let text = unsafe { std::ffi::CStr::from_ptr(qz_last_global_error()) }.to_string_lossy().into_owned();
            return Err(EncodingError::EncodingError(text));
        }
        Ok(Self(ctx))
    }
}
impl Drop for CudaContext {
    fn drop(&mut self) {
        unsafe { qz_ctx_destroy(self.0) } // the context lives until its last index is gone (reference counted)
    }
}

/// { context, index }: Drop -> qz_index_destroy.  The host `TStorage` keeps owning the encoded bytes, so
/// get_vector_data / save / load of the reference types are untouched.
pub struct CudaIndex {
    pub ctx: Arc<CudaContext>,
    pub index: *mut qz_index,
}
unsafe impl Send for CudaIndex {}
unsafe impl Sync for CudaIndex {} // entry points lock the context internally
impl Drop for CudaIndex {
    fn drop(&mut self) {
        unsafe { qz_index_destroy(self.index) }
    }
}

impl CudaIndex {
    pub fn check(&self, st: qz_status) -> Result<(), EncodingError> {
        check(self.ctx.0, st)
    }
    /// the encoded query as the byte blob the C-ABI takes (layouts in include/qz_cuda.h)
    pub fn encode_query_blob(&self, query: &[f32]) -> Vec<u8> {
        let mut blob = vec![0u8; unsafe { qz_query_bytes(self.index) } as usize];
        let st = unsafe { qz_encode_query(self.index, query.as_ptr(), blob.as_mut_ptr() as *mut _) };
        debug_assert_eq!(st, 0);
        blob
    }
    pub fn score_points(&self, blob: &[u8], ids: &[u32]) -> Vec<f32> {
        let mut out = vec![0f32; ids.len()];
        unsafe { qz_score_points(self.index, blob.as_ptr() as *const _, ids.as_ptr(), ids.len() as u64, out.as_mut_ptr()) };
        out
    }
    pub fn score_pairs(&self, a: &[u32], b: &[u32]) -> Vec<f32> {
        let mut out = vec![0f32; a.len()];
        unsafe { qz_score_pairs(self.index, a.as_ptr(), b.as_ptr(), a.len() as u64, out.as_mut_ptr()) };
        out
    }
    pub fn score_all(&self, blob: &[u8]) -> Vec<f32> {
        let mut out = vec![0f32; unsafe { qz_index_count(self.index) } as usize];
        unsafe { qz_scan_scores(self.index, blob.as_ptr() as *const _, out.as_mut_ptr()) };
        out
    }
    /// best first; ties to the lower id (the ordering contract of include/qz_cuda.h)
    pub fn search_top_k(&self, blob: &[u8], k: usize, largest: bool) -> Vec<(f32, u32)> {
        let (mut s, mut i, mut n) = (vec![0f32; k], vec![0u32; k], 0u32);
        unsafe { qz_scan_topk(self.index, blob.as_ptr() as *const _, k as u32, largest as i32, s.as_mut_ptr(), i.as_mut_ptr(), &mut n) };
        s.into_iter().zip(i).take(n as usize).collect()
    }
    /// raw f32 query: encode_query + scan + top-k in one call and one synchronisation
    pub fn search(&self, query: &[f32], k: usize, largest: bool) -> Vec<(f32, u32)> {
        let (mut s, mut i, mut n) = (vec![0f32; k], vec![0u32; k], 0u32);
        unsafe { qz_search_topk(self.index, query.as_ptr(), k as u32, largest as i32, s.as_mut_ptr(), i.as_mut_ptr(), &mut n) };
        s.into_iter().zip(i).take(n as usize).collect()
    }
}

/// The reference polls `stop_condition()` between vectors (encoded_vectors_u8.rs:74-76, binary.rs:178-180,
/// pq.rs:198-200, kmeans.rs:29-31); the library polls an i32 between row pieces / Lloyd iterations.  A watcher thread
/// copies one into the other for the duration of an encode.
pub struct StopFlag {
    flag: Arc<AtomicI32>,
    done: Arc<AtomicI32>,
    watcher: Option<std::thread::JoinHandle<()>>,
}
impl StopFlag {
    pub fn watch(stop_condition: impl Fn() -> bool + Send + 'static) -> Self {
        let (flag, done) = (Arc::new(AtomicI32::new(0)), Arc::new(AtomicI32::new(0)));
        let (f, d) = (flag.clone(), done.clone());
        let watcher = std::thread::spawn(move || {
            while d.load(Ordering::Relaxed) == 0 {
                if stop_condition() {
                    f.store(1, Ordering::Relaxed);
                    return;
                }
                std::thread::sleep(std::time::Duration::from_millis(1));
            }
        });
        Self { flag, done, watcher: Some(watcher) }
    }
    pub fn ptr(&self) -> *const i32 {
        self.flag.as_ptr() as *const i32
    }
}
impl Drop for StopFlag {
    fn drop(&mut self) {
        self.done.store(1, Ordering::Relaxed);
        if let Some(w) = self.watcher.take() {
            let _ = w.join();
        }
    }
}
