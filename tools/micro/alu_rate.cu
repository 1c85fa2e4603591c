// alu_rate.cu — issue-rate microbenchmark of the ops the batched epilogue can be built from (VIADDMNMX, IADD + IMNMX,
// I2F + FFMA + FSETP, IADD + ISETP + mask).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o alu_rate alu_rate.cu
// Eight accumulators per thread, each op reads a neighbour accumulator, so nothing is loop invariant.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 8192;

template <int OP>
__global__ void __launch_bounds__(512) k(const int *in, int *out, int A) {
    int a[8], v[8];
    float f[8];
    for (int i = 0; i < 8; i++) { a[i] = in[threadIdx.x + 32 * i]; v[i] = in[threadIdx.x + 7 * i + 1]; f[i] = __int_as_float(a[i]); }
    unsigned mask = 0;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int n = (i + 1) & 7;
            if (OP == 0) a[i] = __viaddmax_s32(a[n], v[i], a[i]);
            else if (OP == 1) a[i] = __vimax3_s32(a[n], v[i], a[i]);
            else if (OP == 2) { float x = fmaf(1.5f, __int2float_rn(a[n]), f[i]); bool p = x < f[n]; mask |= p ? (1u << i) : 0u; a[i] += p ? 3 : 1; }
            else if (OP == 3) { bool p = a[n] + v[i] >= A; mask |= p ? (1u << i) : 0u; a[i] += p ? 3 : 1; }
            else if (OP == 4) f[i] = fmaf(f[n], 1.0001f, f[i]);
            else if (OP == 5) f[i] += __int2float_rn(a[n]), a[i] += 1;
            else if (OP == 6) a[i] = max(a[n] + v[i], a[i] ^ 1);
            else if (OP == 7) a[i] = a[n] + v[i] + a[i];
        }
    }
    int s = (int)mask;
    for (int i = 0; i < 8; i++) s += a[i] + (int)f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char *name, const int *in, int *out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<148 * 2, 512>>>(in, out, 12345);
    cudaEventRecord(e0);
    k<OP><<<148 * 2, 512>>>(in, out, 12345);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double elems = 148.0 * 2 * 512 * ITERS * 8;
    printf("%-28s %.3f ms  %.1f G elem/s  = %.1f elem/clk/SM @1.9GHz\n", name, ms, elems / ms / 1e6, elems / (ms * 1e-3) / 148 / 1.9e9);
}

int main() {
    int *in, *out;
    cudaMalloc(&in, 1 << 20); cudaMemset(in, 1, 1 << 20);
    cudaMalloc(&out, 148 * 2 * 512 * 4);
    run<0>("VIADDMNMX", in, out);
    run<1>("VIMNMX3", in, out);
    run<2>("I2F+FFMA+FSETP+mask", in, out);
    run<3>("IADD+ISETP+mask", in, out);
    run<4>("FFMA", in, out);
    run<5>("I2F+FADD+IADD", in, out);
    run<6>("IADD+LOP+IMNMX", in, out);
    run<7>("IADD3", in, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
