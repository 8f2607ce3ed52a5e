// Microbenchmark: how fast can ONE lane stream L2-resident data into shared memory with 1-D cp.async.bulk copies?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_rate bulk_rate.cu && ./bulk_rate
// Prints bytes / cycle / SM for (bytes per copy) x (copies in flight) x (CTAs).  Answers whether the weight ring of the
// chain engine (drm_tc_chain.cuh: 32 KB slabs, 2-4 in flight) is bound by the copy engine or by its consumers.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__global__ void __launch_bounds__(64, 1) stream_kernel(const uint8_t* src, size_t footprint, uint32_t bytes, int depth, int iters, int shared_src,
                                                         long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bars[8];
    if (threadIdx.x == 0) {
        for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint8_t* base = src + (shared_src ? 0 : (size_t)blockIdx.x * 65536 % footprint);
        size_t off = 0;
        const long long t0 = clock64();
        // prologue: fill the pipeline
        for (int i = 0; i < depth; ++i) {
            mbar_expect_tx(&bars[i], bytes);
            bulk_g2s(smem + (size_t)i * bytes, base + off, bytes, &bars[i]);
            off = (off + bytes) % (footprint - 65536 - bytes);
        }
        uint32_t phase = 0;
        int st = 0;
        for (int it = 0; it < iters; ++it) {
            mbar_wait(&bars[st], phase);
            mbar_expect_tx(&bars[st], bytes);
            bulk_g2s(smem + (size_t)st * bytes, base + off, bytes, &bars[st]);
            off = (off + bytes) % (footprint - 65536 - bytes);
            if (++st == depth) { st = 0; phase ^= 1; }
        }
        for (int i = 0; i < depth; ++i) {
            mbar_wait(&bars[st], phase);
            if (++st == depth) { st = 0; phase ^= 1; }
        }
        cycles[blockIdx.x] = clock64() - t0;
    }
}

int main() {
    const size_t footprint = 4u << 20;  // 4 MB: L2-resident
    uint8_t* src;
    long long* cyc;
    cudaMalloc(&src, footprint);
    cudaMemset(src, 1, footprint);
    cudaMalloc(&cyc, 148 * sizeof(long long));
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 400;
    printf("%8s %6s %6s %7s %12s\n", "bytes", "depth", "ctas", "shared", "B/cycle/SM");
    for (int shared_src = 1; shared_src >= 0; --shared_src)
        for (int ctas : {1, 40, 148})
            for (uint32_t bytes : {4096u, 8192u, 16384u, 32768u, 49152u})
                for (int depth : {1, 2, 4, 8}) {
                    if ((size_t)bytes * depth > 196 * 1024) continue;
                    for (int rep = 0; rep < 2; ++rep) stream_kernel<<<ctas, 64, 200 * 1024>>>(src, footprint, bytes, depth, iters, shared_src, cyc);
                    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
                    long long h[148];
                    cudaMemcpy(h, cyc, ctas * sizeof(long long), cudaMemcpyDeviceToHost);
                    long long mx = 0;
                    for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
                    printf("%8u %6d %6d %7d %12.1f\n", bytes, depth, ctas, shared_src, (double)bytes * (iters + depth) / (double)mx);
                }
    return 0;
}
