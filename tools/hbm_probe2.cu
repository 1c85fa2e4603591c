// hbm_probe2.cu — what slows the bulk-copy ring when real consumers are attached? (experiment only)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../quantization_b200/csrc/common.cuh"
using namespace qz;

// mode bit0: all lanes poll (else lane 0 polls + syncwarp); bit1: read the tile with LDS; bit2: nanosleep backoff
__global__ void __launch_bounds__(544) ring2(const uint8_t *src, size_t total, uint32_t chunk, uint32_t n_slots,
                                             uint32_t n_warps, int mode, uint32_t *out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)n_slots * chunk);
    uint64_t *empty = full + 32;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t wps = n_warps / n_slots;
    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < n_slots; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], wps); }
        fence_mbar_init();
    }
    __syncthreads();
    const uint32_t n_chunks = total / chunk;
    const uint32_t mine = (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
    if (warp == n_warps) {
        if (lane == 0) {
            uint64_t pol = l2_policy_evict_first();
            uint32_t slot = 0, round = 0;
            for (uint32_t t = 0; t < mine; t++) {
                if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
                size_t off = (blockIdx.x + (size_t)t * gridDim.x) * (size_t)chunk;
                mbar_arrive_expect_tx(&full[slot], chunk);
                bulk_g2s_hint(smem + (size_t)slot * chunk, src + off, chunk, &full[slot], pol);
                if (++slot == n_slots) { slot = 0; round++; }
            }
        }
        return;
    }
    if (warp > n_warps) return;
    const uint32_t slot = warp / wps, member = warp % wps;
    const uint32_t *sl = reinterpret_cast<const uint32_t *>(smem + (size_t)slot * chunk);
    const uint32_t words = chunk / 4, per = words / wps;
    uint32_t acc = 0, parity = 0;
    for (uint32_t t = slot; t < mine; t += n_slots, parity ^= 1) {
        if (mode & 1) {
            if (mode & 4) { while (!mbar_try_wait(&full[slot], parity)) __nanosleep(200); }
            else mbar_wait(&full[slot], parity);
        } else {
            if (lane == 0) {
                if (mode & 4) { while (!mbar_try_wait(&full[slot], parity)) __nanosleep(200); }
                else mbar_wait(&full[slot], parity);
            }
            __syncwarp();
        }
        if (mode & 2) {
            const uint32_t *b = sl + member * per;
#pragma unroll 8
            for (uint32_t i = lane; i < per; i += 32) acc = __dp4a(b[i], 0x01020304u, acc);
        } else acc ^= sl[0];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }
    if (acc == 0x12345678u) out[0] = acc;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

int main() {
    const size_t total = (size_t)1540 * 1000000 / 16 * 16;
    uint8_t *d; uint32_t *out;
    CK(cudaMalloc(&d, total + 65536)); CK(cudaMalloc(&out, 64));
    // random-ish content
    { uint32_t *h = (uint32_t *)malloc(total + 65536); uint32_t x = 12345; for (size_t i = 0; i < (total + 65536) / 4; i++) { x = x * 1664525u + 1013904223u; h[i] = x & 0x7f7f7f7f; }
      CK(cudaMemcpy(d, h, total + 65536, cudaMemcpyHostToDevice)); free(h); }
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int reps = 30;
    CK(cudaFuncSetAttribute(ring2, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    struct Cfg { uint32_t chunk, slots, warps; int mode; };
    Cfg cfgs[] = {
        {24704, 8, 8, 0}, {24704, 8, 8, 1}, {24704, 8, 8, 2}, {24704, 8, 8, 3}, {24704, 8, 8, 7}, {24704, 8, 8, 6},
        {24704, 8, 16, 0}, {24704, 8, 16, 1}, {24704, 8, 16, 2}, {24704, 8, 16, 3}, {24704, 8, 16, 7}, {24704, 8, 16, 6},
        {24704, 4, 16, 0}, {24704, 4, 16, 1}, {24704, 4, 16, 2}, {24704, 4, 16, 3}, {24704, 4, 16, 6},
        {24704, 4, 4, 0}, {24704, 4, 4, 2}, {24704, 4, 8, 2}, {24704, 8, 1, 0},
        {49408, 4, 16, 2}, {49408, 4, 16, 3}, {49408, 2, 16, 2}, {12352, 8, 16, 2}, {12352, 16, 16, 2},
    };
    for (auto c : cfgs) {
        size_t smem = (size_t)c.chunk * c.slots + 1024;
        for (int r = -3; r < reps; r++) {
            if (r == 0) cudaEventRecord(e0);
            ring2<<<148, (c.warps + 1) * 32, smem>>>(d, total, c.chunk, c.slots, c.warps, c.mode, out);
        }
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("chunk=%u slots=%u warps=%u mode=%d(%s%s%s)  %8.4f ms  %8.1f GB/s\n", c.chunk, c.slots, c.warps, c.mode,
               (c.mode & 1) ? "all-lanes-poll" : "lane0-poll", (c.mode & 2) ? "+read" : "", (c.mode & 4) ? "+sleep" : "",
               ms / reps, total / (ms / reps * 1e-3) / 1e9);
        fflush(stdout);
    }
    CK(cudaGetLastError());
    return 0;
}
