// Microbenchmark: TMEM read rate seen by the epilogue warps (tcgen05.ld 32x32b.x8 / .x16), W warps per CTA, one CTA per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I dreamer_pytorch_b200/csrc -o /tmp/tmem_rate tools/micro/tmem_rate.cu
// Prints bytes / cycle / SM.  Answers whether the chain engine's epilogues (16 warps, two x8 loads per wait) are bound by
// TMEM read bandwidth.
#include <cstdio>
#include "drm_tc.cuh"

using namespace drm::tc;

template <int WIDE>
__global__ void __launch_bounds__(512, 1) tmem_read_kernel(int iters, int per_wait, long long* cycles, float* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc(&slot, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        float v[4][16];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k < per_wait) {
                const uint32_t col = (uint32_t)(((warp >> 2) * 128 + ((it * 4 + k) * WIDE) % 128));
                if (WIDE == 8) tmem_ld8(base + col, reinterpret_cast<float(&)[8]>(v[k]));
                else tmem_ld16(base + col, v[k]);
            }
        }
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k < per_wait) acc += v[k][0] + v[k][WIDE - 1];
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    if (acc == 123.456f) sink[0] = acc;
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(slot, 512);
}

int main() {
    long long* cyc;
    float* sink;
    cudaMalloc(&cyc, 148 * sizeof(long long));
    cudaMalloc(&sink, 4);
    const int iters = 2000;
    printf("%6s %6s %9s %14s\n", "width", "warps", "per_wait", "B/cycle/SM");
    for (int wide : {8, 16})
        for (int warps : {4, 8, 16})
            for (int per_wait : {1, 2, 4}) {
                for (int rep = 0; rep < 2; ++rep) {
                    if (wide == 8) tmem_read_kernel<8><<<148, warps * 32, 0>>>(iters, per_wait, cyc, sink);
                    else tmem_read_kernel<16><<<148, warps * 32, 0>>>(iters, per_wait, cyc, sink);
                }
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                long long h[148], mx = 0;
                cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
                for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
                printf("%6d %6d %9d %14.1f\n", wide, warps, per_wait, (double)warps * 32 * wide * 4 * per_wait * iters / (double)mx);
            }
    return 0;
}
