// Layout probe: which shared-memory word does tcgen05.mma read for element (m, k) of the A operand under a given
// descriptor / major-ness / kind?  A's 64 KB region holds small integers (word index pieces, exact in tf32 and bf16),
// B is a K-major identity, so one MMA returns D[m][n] = A_as_read_by_the_hardware[m][n] for n < K.
// Used to pin down the MN-major tf32 operand layout of the weight-gradient kernel; not on any hot path.
#include "common.cuh"
#include "dtf_b200.h"

namespace dtf {

// kind 0: bf16 (K = 16 per MMA), kind 1: tf32 (K = 8 per MMA)
__global__ void __launch_bounds__(128, 1)
probe_umma_kernel(int kind, int a_mn, uint32_t lbo, uint32_t sbo, int layout, int pass, float* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int K = kind == 1 ? 8 : 16;
    const int N = 16;
    uint8_t* a_smem = smem;                    // 64 KB
    uint8_t* b_smem = smem + 64 * 1024;        // K-major identity [N][K]
    // A: word w (32-bit) of the region encodes its own index in two passes of 10 bits
    for (int w = threadIdx.x; w < 64 * 1024 / 4; w += blockDim.x) {
        const float v = static_cast<float>(pass == 0 ? (w & 1023) : (w >> 10));
        if (kind == 1) {
            reinterpret_cast<float*>(a_smem)[w] = v;
        } else {                                // bf16: each 16-bit half carries the half-word index piece
            const int h0 = 2 * w, h1 = 2 * w + 1;
            const float v0 = static_cast<float>(pass == 0 ? (h0 & 127) : (h0 >> 7) & 127);
            const float v1 = static_cast<float>(pass == 0 ? (h1 & 127) : (h1 >> 7) & 127);
            reinterpret_cast<uint32_t*>(a_smem)[w] = pack2(v0, v1);
        }
    }
    // B[n][k] = (n == k), K-major SWIZZLE_NONE: core matrix = 8 rows (n) x 16 bytes; LBO (next 16 bytes of K) = 256,
    // SBO (next 8 rows) = 128
    for (int i = threadIdx.x; i < 4096 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(b_smem)[i] = 0u;
    __syncthreads();
    if (threadIdx.x < N) {
        const int n = threadIdx.x;
        if (n < K) {
            const int k = n;
            const int esz = kind == 1 ? 4 : 2;
            const int per16 = 16 / esz;
            const uint32_t off = (n / 8) * 128 + (k / per16) * 256 + (n % 8) * 16 + (k % per16) * esz;
            if (kind == 1) *reinterpret_cast<float*>(b_smem + off) = 1.f;
            else *reinterpret_cast<uint16_t*>(b_smem + off) = 0x3f80;       // bf16 1.0
        }
    }
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 32);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = tmem_base_s;
    if (warp == 1) {
        const uint32_t idesc = kind == 1 ? umma_idesc_tf32(128, N, a_mn != 0, false) : umma_idesc_bf16(128, N, a_mn != 0, false);
        // layout type (descriptor bits 61..63): 0 none, 1 128B with 32-byte base, 2 128B, 4 64B, 6 32B
        const uint64_t adesc = umma_desc(smem_u32(a_smem), lbo, sbo) | (static_cast<uint64_t>(layout & 7) << 61);
        const uint64_t bdesc = umma_desc(smem_u32(b_smem), 256, 128);
        if (elect_one()) {
            if (kind == 1) umma_tf32(tmem_base, adesc, bdesc, idesc, 0);
            else umma_bf16(tmem_base, adesc, bdesc, idesc, 0);
            umma_commit(smem_u32(&bar));
        }
        __syncwarp();
    }
    mbar_wait(smem_u32(&bar), 0, 0x77);
    tc_fence_after_sync();
    float v[16];
    tmem_ld16(tmem_base + ((warp * 32u) << 16), v);
    const int m = warp * 32 + lane;
    for (int n = 0; n < N; ++n) out[m * N + n] = v[n];
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem_base, 32); }
}

}  // namespace dtf

extern "C" int dtf_probe_umma(int kind, int a_mn, unsigned int lbo, unsigned int sbo, int layout, int pass, float* out_dev,
                              cudaStream_t stream) {
    static int set = 0;
    if (!set) {
        if (cudaFuncSetAttribute(dtf::probe_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        set = 1;
    }
    dtf::probe_umma_kernel<<<1, 128, 72 * 1024, stream>>>(kind, a_mn, lbo, sbo, layout, pass, out_dev);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
