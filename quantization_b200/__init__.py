"""quantization_b200 — B200-native (sm_100a) drop-in for the encoded-vector hot path of
qdrant/quantization.  The product is libqz_cuda.so (C-ABI in include/qz_cuda.h); this package is
the host-side mirror of the reference's `EncodedVectors` trait over that library, used by the
parity tests and the benchmark.  There is no CPU fallback: without the built library and a
B200 every call raises."""
from .encoded_vectors import (Context, DistanceType, EncodingError, VecStorageBuilder, VectorParameters,  # noqa: F401
                              default_context, get_vector_data)
from .encoded_vectors_u8 import EncodedVectorsU8  # noqa: F401
from .encoded_vectors_binary import BITS_STORE_U8, BITS_STORE_U128, EncodedVectorsBin  # noqa: F401
from .encoded_vectors_pq import EncodedVectorsPQ  # noqa: F401
