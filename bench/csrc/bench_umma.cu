// Micro-benchmark: cycles per tcgen05.mma for a given operand layout / shape (one CTA).
// Used to choose the shared-memory layout of the convolution kernels; not on any hot path.
#include "common.cuh"
#include "dtf_b200.h"

namespace dtf {

DTF_DEVINL uint64_t umma_desc_sw(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
    uint64_t d = umma_desc(smem_addr, lbo_bytes, sbo_bytes);
    d |= static_cast<uint64_t>((smem_addr >> 7) & 7u) << 49;     // base offset for unaligned starts
    d |= static_cast<uint64_t>(layout & 7u) << 61;
    return d;
}

__global__ void __launch_bounds__(128, 1)
bench_umma_kernel(int layout, int n, int iters, int a_shift_bytes, int mn_major, unsigned long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = tmem_base_s;
    if (warp == 1) {
        const uint32_t idesc = umma_idesc_bf16(128, n, mn_major, mn_major);
        const uint32_t a_base = smem_u32(smem) + a_shift_bytes;
        const uint32_t b_base = smem_u32(smem) + 96 * 1024;
        // K-major: SBO = 8 rows * row pitch; row pitch = 16 (none), 32, 64, 128 bytes
        uint32_t pitch = layout == 0 ? 16 : layout == 6 ? 32 : layout == 4 ? 64 : 128;
        uint32_t lbo = layout == 0 ? 2048 : 0, sbo = 8 * pitch;
        if (mn_major) { lbo = 128; sbo = 2048; }
        const uint64_t adesc = umma_desc_sw(a_base, lbo, sbo, layout);
        const uint64_t bdesc = umma_desc_sw(b_base, layout == 0 ? n * 16 : 0, sbo, layout);
        const unsigned long long t0 = clock64();
        for (int i = 0; i < iters; ++i)
            if (elect_one()) umma_bf16(tmem_base, adesc, bdesc, idesc, i != 0);
        if (elect_one()) umma_commit(smem_u32(&bar));
        __syncwarp();
        const unsigned long long t1 = clock64();
        mbar_wait(smem_u32(&bar), 0, 0x99);
        const unsigned long long t2 = clock64();
        if (threadIdx.x == 32) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace dtf

extern "C" int dtf_bench_umma(int layout, int n, int iters, int a_shift_bytes, int mn_major,
                              unsigned long long* out_dev, cudaStream_t stream) {
    static int set = 0;
    if (!set) {
        if (cudaFuncSetAttribute(dtf::bench_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        set = 1;
    }
    dtf::bench_umma_kernel<<<1, 128, 200 * 1024, stream>>>(layout, n, iters, a_shift_bytes, mn_major, out_dev);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
