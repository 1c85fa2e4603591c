// encoded_vectors_binary_cuda.rs — EncodedVectorsBin behind libqz_cuda (cargo feature `cuda`).  SYNTAX-REVIEWED ONLY.
// Reference: quantization/src/encoded_vectors_binary.rs (types :11-19, BitsStoreType :21-163, encode :165-208,
// calculate_metric :237-252, trait impl :254-314).
use crate::cuda_ffi::*;
use crate::cuda_index::{CudaContext, CudaIndex, StopFlag};
use crate::encoded_storage::{EncodedStorage, EncodedStorageBuilder};
use crate::encoded_vectors::{EncodedVectors, VectorParameters};
use crate::EncodingError;
use std::marker::PhantomData;
use std::path::Path;
use std::sync::Arc;

pub struct EncodedVectorsBin<TBitsStoreType: BitsStoreType, TStorage: EncodedStorage> {
    encoded_vectors: TStorage,
    metadata: Metadata, // { vector_parameters }, binary.rs:21-24
    dev: CudaIndex,
    bits_store_type: PhantomData<TBitsStoreType>,
}

pub struct EncodedBinVector<TBitsStoreType: BitsStoreType> {
    encoded_vector: Vec<TBitsStoreType>, // binary.rs:17-19; its bytes ARE the blob the C-ABI takes
}

impl<TBitsStoreType: BitsStoreType, TStorage: EncodedStorage> EncodedVectorsBin<TBitsStoreType, TStorage> {
    pub fn encode<'a>(
        orig_data: impl Iterator<Item = impl AsRef<[f32]> + 'a> + Clone,
        mut storage_builder: impl EncodedStorageBuilder<TStorage>,
        vector_parameters: &VectorParameters,
        stop_condition: impl Fn() -> bool + Send + 'static,
    ) -> Result<Self, EncodingError> {
        let ctx = Arc::new(CudaContext::new(0)?);
        let store = std::mem::size_of::<TBitsStoreType>() as i32; // u8 -> 1, u128 -> 16 (binary.rs:26-163)
        let (count, dim) = (vector_parameters.count, vector_parameters.dim);
        let flat: Vec<f32> = orig_data.flat_map(|v| v.as_ref().to_vec()).collect();
        let row = unsafe { qz_bin_row_bytes(dim as u32, store) } as usize; // get_quantized_vector_size_from_params, :210-212
        let mut bytes = vec![0u8; row * count];
        let stop = StopFlag::watch(stop_condition); // binary.rs:178-180
        check(ctx.0, unsafe { qz_bin_encode(ctx.0, flat.as_ptr(), count as u64, dim as u32, store, stop.ptr(), bytes.as_mut_ptr()) })?;
        drop(stop);
        for r in bytes.chunks_exact(row) {
            storage_builder.push_vector_data(r);
        }
        Self::attach(storage_builder.build(), Metadata { vector_parameters: vector_parameters.clone() }, ctx, &bytes)
    }

    fn attach(encoded_vectors: TStorage, m: Metadata, ctx: Arc<CudaContext>, bytes: &[u8]) -> Result<Self, EncodingError> {
        let p = &m.vector_parameters;
        let mut index = std::ptr::null_mut();
        check(ctx.0, unsafe {
            qz_bin_index_create(ctx.0, bytes.as_ptr(), p.count as u64, p.dim as u32, std::mem::size_of::<TBitsStoreType>() as i32,
                                p.distance_type as i32, p.invert as i32, 0, &mut index)
        })?;
        Ok(Self { encoded_vectors, metadata: m, dev: CudaIndex { ctx, index }, bits_store_type: PhantomData })
    }

    fn blob(q: &EncodedBinVector<TBitsStoreType>) -> &[u8] {
        unsafe { std::slice::from_raw_parts(q.encoded_vector.as_ptr() as *const u8, std::mem::size_of_val(&q.encoded_vector[..])) }
    }

    pub fn score_all(&self, q: &EncodedBinVector<TBitsStoreType>) -> Vec<f32> { self.dev.score_all(Self::blob(q)) }
    pub fn search_top_k(&self, q: &EncodedBinVector<TBitsStoreType>, k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search_top_k(Self::blob(q), k, largest) }
    pub fn score_points(&self, q: &EncodedBinVector<TBitsStoreType>, ids: &[u32]) -> Vec<f32> { self.dev.score_points(Self::blob(q), ids) }
    pub fn search(&self, query: &[f32], k: usize, largest: bool) -> Vec<(f32, u32)> { self.dev.search(query, k, largest) }
}

impl<TBitsStoreType: BitsStoreType, TStorage: EncodedStorage> EncodedVectors<EncodedBinVector<TBitsStoreType>>
    for EncodedVectorsBin<TBitsStoreType, TStorage>
{
    fn save(&self, data_path: &Path, meta_path: &Path) -> std::io::Result<()> {
        std::fs::write(meta_path, serde_json::to_string(&self.metadata)?)?; // binary.rs:260-270
        self.encoded_vectors.save_to_file(data_path)
    }

    fn load(data_path: &Path, meta_path: &Path, vector_parameters: &VectorParameters) -> std::io::Result<Self> {
        let metadata: Metadata = serde_json::from_str(&std::fs::read_to_string(meta_path)?)?; // binary.rs:272-286
        let row = unsafe { qz_bin_row_bytes(vector_parameters.dim as u32, std::mem::size_of::<TBitsStoreType>() as i32) } as usize;
        let encoded_vectors = TStorage::from_file(data_path, row, vector_parameters.count)?;
        let bytes: Vec<u8> = (0..vector_parameters.count).flat_map(|i| encoded_vectors.get_vector_data(i, row).to_vec()).collect();
        let err = |e: EncodingError| std::io::Error::new(std::io::ErrorKind::Other, format!("{e:?}"));
        Self::attach(encoded_vectors, metadata, Arc::new(CudaContext::new(0).map_err(err)?), &bytes).map_err(err)
    }

    fn encode_query(&self, query: &[f32]) -> EncodedBinVector<TBitsStoreType> {
        // binary.rs:288-291 (= encode_vector :214-235: bit j set iff query[j] > 0) on the device
        let blob = self.dev.encode_query_blob(query);
        let n = blob.len() / std::mem::size_of::<TBitsStoreType>();
        let mut encoded_vector = Vec::<TBitsStoreType>::with_capacity(n);
        unsafe {
            std::ptr::copy_nonoverlapping(blob.as_ptr(), encoded_vector.as_mut_ptr() as *mut u8, blob.len());
            encoded_vector.set_len(n);
        }
        EncodedBinVector { encoded_vector }
    }

    fn score_point(&self, query: &EncodedBinVector<TBitsStoreType>, i: u32) -> f32 {
        self.dev.score_points(Self::blob(query), &[i])[0] // binary.rs:293-300; batch ids where possible
    }

    fn score_internal(&self, i: u32, j: u32) -> f32 {
        self.dev.score_pairs(&[i], &[j])[0] // binary.rs:302-314
    }
}
