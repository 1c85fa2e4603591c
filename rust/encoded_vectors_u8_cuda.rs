// encoded_vectors_u8_cuda.rs — EncodedVectorsU8 behind libqz_cuda (cargo feature `cuda`).  SYNTAX-REVIEWED ONLY.
// Reference: quantization/src/encoded_vectors_u8.rs (struct :14-31, encode :34-140, trait impl :262-453).
// The type keeps the reference's fields and file formats; only the scorers move to the device.
use crate::cuda_ffi::*;
use crate::cuda_index::{CudaContext, CudaIndex, StopFlag};
use crate::encoded_storage::{EncodedStorage, EncodedStorageBuilder};
use crate::encoded_vectors::{EncodedVectors, VectorParameters};
use crate::EncodingError;
use std::path::Path;
use std::sync::Arc;

pub struct EncodedVectorsU8<TStorage: EncodedStorage> {
    encoded_vectors: TStorage, // host bytes: get_vector_data / save keep working (u8.rs:263-272)
    metadata: Metadata,        // { actual_dim, alpha, offset, multiplier, vector_parameters }, u8.rs:24-31, unchanged
    dev: CudaIndex,
}

pub struct EncodedQueryU8 {
    offset: f32,
    encoded_query: Vec<u8>,
    blob: Vec<u8>, // [f32 offset][codes]: the form the C-ABI takes, kept so score_point does not rebuild it
}

impl<TStorage: EncodedStorage> EncodedVectorsU8<TStorage> {
    pub fn encode<'a>(
        orig_data: impl Iterator<Item = impl AsRef<[f32]> + 'a> + Clone,
        mut storage_builder: impl EncodedStorageBuilder<TStorage>,
        vector_parameters: &VectorParameters,
        quantile: Option<f32>,
        stop_condition: impl Fn() -> bool + Send + 'static,
    ) -> Result<Self, EncodingError> {
        crate::encoded_vectors::validate_vector_parameters(orig_data.clone(), vector_parameters)?; // u8.rs:41
        let ctx = Arc::new(CudaContext::new(0)?);
        let (count, dim) = (vector_parameters.count, vector_parameters.dim);
        // the reference walks the iterator twice (min/max, then codes); the library streams one flat buffer through
        // the device in 256 MB pieces and polls the stop flag between them (u8.rs:74-76)
        let flat: Vec<f32> = orig_data.flat_map(|v| v.as_ref().to_vec()).collect();
        let row = unsafe { qz_u8_row_bytes(dim as u32) } as usize;
        let mut bytes = vec![0u8; row * count];
        let (mut alpha, mut offset, mut multiplier) = (0f32, 0f32, 0f32);
        let stop = StopFlag::watch(stop_condition);
        let q = quantile.as_ref().map_or(std::ptr::null(), |q| q as *const f32);
        check(ctx.0, unsafe {
            qz_u8_encode(ctx.0, flat.as_ptr(), count as u64, dim as u32, vector_parameters.distance_type as i32,
                         vector_parameters.invert as i32, q, std::ptr::null(), 0, rand::random(), stop.ptr(),
                         bytes.as_mut_ptr(), &mut alpha, &mut offset, &mut multiplier)
        })?;
        drop(stop);
        for r in bytes.chunks_exact(row) {
            storage_builder.push_vector_data(r); // EncodedStorageBuilder contract, encoded_storage.rs:27-31
        }
        let metadata = Metadata { actual_dim: unsafe { qz_u8_actual_dim(dim as u32) } as usize, alpha, offset, multiplier,
                                  vector_parameters: vector_parameters.clone() };
        Self::attach(storage_builder.build(), metadata, ctx, &bytes)
    }

    fn attach(encoded_vectors: TStorage, m: Metadata, ctx: Arc<CudaContext>, bytes: &[u8]) -> Result<Self, EncodingError> {
        let p = &m.vector_parameters;
        let mut index = std::ptr::null_mut();
        check(ctx.0, unsafe {
            qz_u8_index_create(ctx.0, bytes.as_ptr(), p.count as u64, p.dim as u32, p.distance_type as i32,
                               p.invert as i32, m.alpha, m.offset, m.multiplier, 0, &mut index)
        })?;
        Ok(Self { encoded_vectors, metadata: m, dev: CudaIndex { ctx, index } })
    }

    // ---- what makes the GPU path fast: whole-corpus calls instead of the caller's loop + BinaryHeap
    //      (demos/src/ann_benchmark_data.rs:150-167)
    pub fn score_all(&self, q: &EncodedQueryU8) -> Vec<f32> { self.dev.score_all(&q.blob) }
    pub fn search_top_k(&self, q: &EncodedQueryU8, k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search_top_k(&q.blob, k, largest) }
    pub fn score_points(&self, q: &EncodedQueryU8, ids: &[u32]) -> Vec<f32> { self.dev.score_points(&q.blob, ids) }
    pub fn search(&self, query: &[f32], k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search(query, k, largest) }
    /// nq raw queries at once (Dot): int8 tensor-core contraction + per-query top-k; [nq][k]
    pub fn search_batch(&self, queries: &[f32], nq: usize, k: usize, largest: bool) -> Result<(Vec<f32>, Vec<u32>), EncodingError> {
        let (mut s, mut i) = (vec![0f32; nq * k], vec![0u32; nq * k]);
        self.dev.check(unsafe { qz_u8_search_batch(self.dev.index, queries.as_ptr(), nq as u32, k as u32, largest as i32, s.as_mut_ptr(), i.as_mut_ptr()) })?;
        Ok((s, i))
    }
}

impl<TStorage: EncodedStorage> EncodedVectors<EncodedQueryU8> for EncodedVectorsU8<TStorage> {
    fn save(&self, data_path: &Path, meta_path: &Path) -> std::io::Result<()> {
        // unchanged host code, u8.rs:263-272: metadata as JSON, storage bytes through TStorage
        std::fs::write(meta_path, serde_json::to_string(&self.metadata)?)?;
        self.encoded_vectors.save_to_file(data_path)
    }

    fn load(data_path: &Path, meta_path: &Path, vector_parameters: &VectorParameters) -> std::io::Result<Self> {
        // u8.rs:274-288, then one upload
        let metadata: Metadata = serde_json::from_str(&std::fs::read_to_string(meta_path)?)?;
        let row = unsafe { qz_u8_row_bytes(vector_parameters.dim as u32) } as usize;
        let encoded_vectors = TStorage::from_file(data_path, row, vector_parameters.count)?;
        let bytes: Vec<u8> = (0..vector_parameters.count).flat_map(|i| encoded_vectors.get_vector_data(i, row).to_vec()).collect();
        let ctx = Arc::new(CudaContext::new(0).map_err(|e| std::io::Error::new(std::io::ErrorKind::Other, format!("{e:?}")))?);
        Self::attach(encoded_vectors, metadata, ctx, &bytes).map_err(|e| std::io::Error::new(std::io::ErrorKind::Other, format!("{e:?}")))
    }

    fn encode_query(&self, query: &[f32]) -> EncodedQueryU8 {
        // u8.rs:290-329 on the device; blob = [f32 offset][actual_dim codes]
        let blob = self.dev.encode_query_blob(query);
        EncodedQueryU8 { offset: f32::from_ne_bytes(blob[..4].try_into().unwrap()), encoded_query: blob[4..].to_vec(), blob }
    }

    fn score_point(&self, query: &EncodedQueryU8, i: u32) -> f32 {
        // u8.rs:331-384.  Correct, but one launch and one 4-byte copy per call (about 10 us): graph-search callers
        // should hand over id lists (score_points) or use search_top_k
        self.dev.score_points(&query.blob, &[i])[0]
    }

    fn score_internal(&self, i: u32, j: u32) -> f32 {
        self.dev.score_pairs(&[i], &[j])[0] // u8.rs:386-453
    }
}
