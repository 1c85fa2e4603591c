// hbm_probe.cu — read-bandwidth ceiling probe for the scan design (not part of the library).
//   A: LDG.128 grid-stride sum;  B: 1-D bulk-copy (TMA) ring with a no-op consumer.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hbm_probe hbm_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../quantization_b200/csrc/common.cuh"
using namespace qz;

template <int U>
__global__ void ldg_kernel(const uint4 *p, size_t n16, uint32_t *out) {
    uint32_t acc = 0;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n16; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = __ldcs(p + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; u++) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) out[0] = acc;
}

__global__ void __launch_bounds__(64) ring_kernel(const uint8_t *src, size_t total, uint32_t chunk, uint32_t n_slots,
                                                  int hint, uint32_t *out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)n_slots * chunk);
    uint64_t *empty = full + 32;
    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < n_slots; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    const size_t n_chunks = total / chunk;
    const size_t mine = (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
    if (threadIdx.x == 0) {
        uint64_t pol = l2_policy_evict_first();
        for (size_t t = 0; t < mine; t++) {
            uint32_t slot = t % n_slots, round = t / n_slots;
            if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
            size_t off = (blockIdx.x + t * gridDim.x) * (size_t)chunk;
            mbar_arrive_expect_tx(&full[slot], chunk);
            if (hint) bulk_g2s_hint(smem + (size_t)slot * chunk, src + off, chunk, &full[slot], pol);
            else bulk_g2s(smem + (size_t)slot * chunk, src + off, chunk, &full[slot]);
        }
    } else if (threadIdx.x == 32) {
        uint32_t acc = 0;
        for (size_t t = 0; t < mine; t++) {
            uint32_t slot = t % n_slots, round = t / n_slots;
            mbar_wait(&full[slot], round & 1);
            acc ^= *reinterpret_cast<uint32_t *>(smem + (size_t)slot * chunk);
            mbar_arrive(&empty[slot]);
        }
        if (acc == 0x12345678u) out[0] = acc;
    }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

int main() {
    const size_t total = (size_t)1540 * 1000000 / 16 * 16;
    uint8_t *d; uint32_t *out;
    CK(cudaMalloc(&d, total + 65536)); CK(cudaMalloc(&out, 64));
    CK(cudaMemset(d, 1, total + 65536));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int reps = 30;
    auto report = [&](const char *name, float ms) { printf("%-48s %8.4f ms  %8.1f GB/s\n", name, ms / reps, total / (ms / reps * 1e-3) / 1e9); fflush(stdout); };
    char name[128];
    // A
    for (int cps : {2, 4, 8}) for (int thr : {256, 512}) {
        for (int u : {4, 8}) {
            for (int r = -3; r < reps; r++) {
                if (r == 0) cudaEventRecord(e0);
                if (u == 4) ldg_kernel<4><<<148 * cps, thr>>>((const uint4 *)d, total / 16, out);
                else ldg_kernel<8><<<148 * cps, thr>>>((const uint4 *)d, total / 16, out);
            }
            cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            snprintf(name, sizeof name, "ldg128 ctas/sm=%d thr=%d unroll=%d", cps, thr, u); report(name, ms);
        }
    }
    // B
    CK(cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    struct Cfg { uint32_t chunk, slots; int cps; int hint; };
    Cfg cfgs[] = {{24704, 8, 1, 1}, {24704, 8, 1, 0}, {24704, 4, 2, 1}, {12352, 16, 1, 1}, {12352, 8, 2, 1}, {6160, 32, 1, 1}, {6160, 16, 2, 1},
                  {6160, 8, 4, 1}, {32768, 6, 1, 1}, {16384, 12, 1, 1}, {16384, 6, 2, 1}, {8192, 24, 1, 1}, {8192, 12, 2, 1}, {4096, 24, 2, 1}, {4096, 12, 4, 1},
                  {24704, 4, 1, 1}, {24704, 2, 1, 1}, {24704, 3, 2, 1}, {49408, 4, 1, 1}, {2048, 24, 4, 1}};
    for (auto c : cfgs) {
        size_t smem = (size_t)c.chunk * c.slots + 1024;
        for (int r = -3; r < reps; r++) {
            if (r == 0) cudaEventRecord(e0);
            ring_kernel<<<148 * c.cps, 64, smem>>>(d, total, c.chunk, c.slots, c.hint, out);
        }
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        snprintf(name, sizeof name, "tma ring chunk=%u slots=%u ctas/sm=%d hint=%d", c.chunk, c.slots, c.cps, c.hint); report(name, ms);
    }
    CK(cudaGetLastError());
    return 0;
}
