// encoded_vectors_pq_cuda.rs — EncodedVectorsPQ behind libqz_cuda (cargo feature `cuda`).  SYNTAX-REVIEWED ONLY.
// Reference: quantization/src/encoded_vectors_pq.rs (types :27-44, encode :56-107, k-means :216-403,
// score_point :405-494, trait impl :496-593).
use crate::cuda_ffi::*;
use crate::cuda_index::{CudaContext, CudaIndex, StopFlag};
use crate::encoded_storage::{EncodedStorage, EncodedStorageBuilder};
use crate::encoded_vectors::{EncodedVectors, VectorParameters};
use crate::EncodingError;
use std::ops::Range;
use std::path::Path;
use std::sync::Arc;

pub const KMEANS_SAMPLE_SIZE: usize = 10_000; // pq.rs:22
pub const KMEANS_MAX_ITERATIONS: usize = 100; // pq.rs:23
pub const KMEANS_ACCURACY: f32 = 1e-5; // pq.rs:24
pub const CENTROIDS_COUNT: usize = 256; // pq.rs:25

pub struct EncodedVectorsPQ<TStorage: EncodedStorage> {
    encoded_vectors: TStorage,
    metadata: Metadata, // { centroids: Vec<Vec<f32>>, vector_division: Vec<Range<usize>>, vector_parameters }, pq.rs:39-44
    dev: CudaIndex,
}

pub struct EncodedQueryPQ {
    lut: Vec<f32>, // pq.rs:35-37: [chunk][256]; its bytes are the blob the C-ABI takes
}

impl<TStorage: EncodedStorage> EncodedVectorsPQ<TStorage> {
    pub fn encode<'a>(
        orig_data: impl Iterator<Item = impl AsRef<[f32]> + 'a> + Clone,
        mut storage_builder: impl EncodedStorageBuilder<TStorage>,
        vector_parameters: &VectorParameters,
        chunk_size: usize,
        _max_kmeans_threads: usize, // the device runs all chunk problems at once (the reference: pq.rs:310-339 serially per chunk)
        stop_condition: impl Fn() -> bool + Send + 'static,
    ) -> Result<Self, EncodingError> {
        let ctx = Arc::new(CudaContext::new(0)?);
        let (count, dim) = (vector_parameters.count, vector_parameters.dim);
        let flat: Vec<f32> = orig_data.flat_map(|v| v.as_ref().to_vec()).collect();
        let vector_division: Vec<Range<usize>> = (0..dim).step_by(chunk_size).map(|a| a..(a + chunk_size).min(dim)).collect(); // pq.rs:116-121
        let mut centroids = vec![0f32; CENTROIDS_COUNT * dim];
        let stop = StopFlag::watch(stop_condition);
        let mut iterations = 0u32;
        // find_centroids (pq.rs:216-308): sample of KMEANS_SAMPLE_SIZE, KMEANS_MAX_ITERATIONS, KMEANS_ACCURACY
        check(ctx.0, unsafe {
            qz_pq_train(ctx.0, flat.as_ptr(), count as u64, dim as u32, chunk_size as u32, std::ptr::null(), 0, rand::random(),
                        KMEANS_MAX_ITERATIONS as u32, KMEANS_ACCURACY, stop.ptr(), centroids.as_mut_ptr(), &mut iterations)
        })?;
        // encode_storage (pq.rs:123-214): nearest centroid per chunk, ties to the lower index
        let mut codes = vec![0u8; count * vector_division.len()];
        check(ctx.0, unsafe {
            qz_pq_assign(ctx.0, flat.as_ptr(), count as u64, dim as u32, chunk_size as u32, centroids.as_ptr(), stop.ptr(), codes.as_mut_ptr())
        })?;
        drop(stop);
        for r in codes.chunks_exact(vector_division.len()) {
            storage_builder.push_vector_data(r);
        }
        let metadata = Metadata { centroids: centroids.chunks_exact(dim).map(|c| c.to_vec()).collect(), vector_division,
                                  vector_parameters: vector_parameters.clone() };
        Self::attach(storage_builder.build(), metadata, ctx, &codes, chunk_size)
    }

    fn attach(encoded_vectors: TStorage, m: Metadata, ctx: Arc<CudaContext>, codes: &[u8], chunk_size: usize) -> Result<Self, EncodingError> {
        let p = &m.vector_parameters;
        let flat_centroids: Vec<f32> = m.centroids.iter().flatten().copied().collect();
        let mut index = std::ptr::null_mut();
        check(ctx.0, unsafe {
            qz_pq_index_create(ctx.0, codes.as_ptr(), p.count as u64, p.dim as u32, chunk_size as u32, flat_centroids.as_ptr(),
                               p.distance_type as i32, p.invert as i32, 0, &mut index)
        })?;
        Ok(Self { encoded_vectors, metadata: m, dev: CudaIndex { ctx, index } })
    }

    fn blob(q: &EncodedQueryPQ) -> &[u8] {
        unsafe { std::slice::from_raw_parts(q.lut.as_ptr() as *const u8, q.lut.len() * 4) }
    }

    /// every score of the corpus: within rtol 1e-5 of score_point's x86 summation order (the scan sums a row's
    /// lookups in its own fixed order); score_points below is bit-exact with score_point_sse
    pub fn score_all(&self, q: &EncodedQueryPQ) -> Vec<f32> { self.dev.score_all(Self::blob(q)) }
    pub fn search_top_k(&self, q: &EncodedQueryPQ, k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search_top_k(Self::blob(q), k, largest) }
    pub fn score_points(&self, q: &EncodedQueryPQ, ids: &[u32]) -> Vec<f32> { self.dev.score_points(Self::blob(q), ids) }
    pub fn search(&self, query: &[f32], k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search(query, k, largest) }
}

impl<TStorage: EncodedStorage> EncodedVectors<EncodedQueryPQ> for EncodedVectorsPQ<TStorage> {
    fn save(&self, data_path: &Path, meta_path: &Path) -> std::io::Result<()> {
        std::fs::write(meta_path, serde_json::to_string(&self.metadata)?)?; // pq.rs:498-508
        self.encoded_vectors.save_to_file(data_path)
    }

    fn load(data_path: &Path, meta_path: &Path, vector_parameters: &VectorParameters) -> std::io::Result<Self> {
        let metadata: Metadata = serde_json::from_str(&std::fs::read_to_string(meta_path)?)?; // pq.rs:510-523
        let row = metadata.vector_division.len();
        let chunk_size = metadata.vector_division.first().map_or(1, |r| r.end - r.start);
        let encoded_vectors = TStorage::from_file(data_path, row, vector_parameters.count)?;
        let codes: Vec<u8> = (0..vector_parameters.count).flat_map(|i| encoded_vectors.get_vector_data(i, row).to_vec()).collect();
        let err = |e: EncodingError| std::io::Error::new(std::io::ErrorKind::Other, format!("{e:?}"));
        Self::attach(encoded_vectors, metadata, Arc::new(CudaContext::new(0).map_err(err)?), &codes, chunk_size).map_err(err)
    }

    fn encode_query(&self, query: &[f32]) -> EncodedQueryPQ {
        // pq.rs:525-547 on the device: lut[chunk * 256 + c] = +-distance(query[range], centroid c[range])
        let blob = self.dev.encode_query_blob(query);
        EncodedQueryPQ { lut: blob.chunks_exact(4).map(|b| f32::from_ne_bytes(b.try_into().unwrap())).collect() }
    }

    fn score_point(&self, query: &EncodedQueryPQ, i: u32) -> f32 {
        self.dev.score_points(Self::blob(query), &[i])[0] // pq.rs:549-564, bit-exact with score_point_sse (:405-440)
    }

    fn score_internal(&self, i: u32, j: u32) -> f32 {
        self.dev.score_pairs(&[i], &[j])[0] // pq.rs:566-593
    }
}
