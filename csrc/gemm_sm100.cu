// Dense-layer GEMM on tcgen05 (MNIST: dense 3136 -> 1024 forward, dgrad and wgrad — reference
// mnist_model.py:90-97 — the only "real" GEMM of the reference) plus the small kernels around it:
// 2x2 max-pool forward / backward (+ReLU mask + bias gradient), activation-mask + bias gradient.
//
// D[M][N] = A x B^T where each operand is a row-major bf16 matrix that is either
//   K-major   X[i][k]  (rows index M or N, columns index K)   or
//   MN-major  X[k][i]  (rows index K, columns index M or N);
// both land in shared memory as 16-byte chunks [chunk][row][8] and are described to the tensor core
// with SWIZZLE_NONE descriptors (K-major: LBO = chunk stride, SBO = 128; MN-major: SBO = chunk stride,
// LBO = 128), exactly like the convolution kernels.  Chunk strides are padded by 16 bytes so that the
// coalesced global->shared gather (consecutive threads = consecutive chunks of one row) is free of bank
// conflicts.  Tile 128 x 128 x 64, three operand stages, two TMEM accumulators are not needed (one tile
// per CTA): 4 epilogue warps + 1 MMA warp + 8 gather warps.
#include <cstdlib>
#include <cuda.h>
#include "common.cuh"
#include "dtf_b200.h"

namespace dtf {

constexpr int kGThreads = 13 * 32;
constexpr int kGGather = 8 * 32;
constexpr int kGStages = 3;
constexpr int kBM = 128, kBN = 128, kBK = 64;
constexpr int kCsK = 128 * 16 + 16;      // K-major chunk stride (128 rows), padded
constexpr int kCsMN = 64 * 16 + 16;      // MN-major chunk stride (64 k-rows), padded
constexpr int kPanelBytes = 16 * kCsMN > 8 * kCsK ? 16 * kCsMN : 8 * kCsK;   // 16640

struct __align__(16) GemmHeader {
    uint64_t full[kGStages], empty[kGStages], acc_full;
    uint32_t tmem_base, pad;
};

DTF_DEVINL uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

// one 128-wide panel (rows0.. of the M or N dimension) for K range [k0, k0+64)
DTF_DEVINL void load_panel(const dtf_bf16* __restrict__ base, long long ld, int mn_major, int rows0, int rows_total,
                           int k0, int K, uint32_t dst, int g) {
    if (!mn_major) {      // X[i][k]: 128 rows x 8 chunks; thread -> (row = g/8 + 32 j, chunk = g%8)
        const int c = g & 7;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = (g >> 3) + 32 * j;
            uint4 v = make_uint4(0, 0, 0, 0);
            const int row = rows0 + r, k = k0 + c * 8;
            if (row < rows_total && k < K) v = ldg16(base + static_cast<long long>(row) * ld + k);
            sts16(dst + c * kCsK + r * 16, v);
        }
    } else {              // X[k][i]: 64 k-rows x 16 chunks; thread -> (k-row = g/16 + 16 j, chunk = g%16)
        const int c = g & 15;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = (g >> 4) + 16 * j;
            uint4 v = make_uint4(0, 0, 0, 0);
            const int k = k0 + r, col = rows0 + c * 8;
            if (k < K && col < rows_total) v = ldg16(base + static_cast<long long>(k) * ld + col);
            sts16(dst + c * kCsMN + r * 16, v);
        }
    }
}

// Epilogue of one 128 x 128 output tile (four warps, one TMEM lane quadrant each): split-K reduction through the
// fp32 workspace when the K loop was split, then bias / ReLU / dropout and the store.  Shared by the gather-fed and
// the TMA-fed kernels.
DTF_DEVINL void gemm_epilogue(const DtfGemm& P, GemmHeader* hdr, uint32_t tmem_base, int tid, int warp, int lane,
                              int m0, int n0, int num_kb) {
    const int m = m0 + warp * 32 + lane;
    const bool mvalid = m < P.M;
    const long long step = P.step ? *P.step : 0;
    bool finish = true;              // this CTA applies the epilogue (always, without split-K)
    if (P.k_split > 1) {
        // 1. add this CTA's partial tile into the fp32 workspace; 2. take a ticket; 3. the last CTA of the
        // tile reads the sum back, leaves zeros behind (the workspace and the counters clean themselves) and
        // runs the normal epilogue.
        if (num_kb > 0) {
            mbar_wait(smem_u32(&hdr->acc_full), 0, 0x53);
            tc_fence_after_sync();
            for (int cg = 0; cg < kBN; cg += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((warp * 32u) << 16) + cg, v);
                const int n = n0 + cg;
                if (!mvalid || n >= P.N) continue;
                float* w = P.ws + static_cast<long long>(m) * P.N + n;
                if (n + 16 <= P.N && (P.N & 3) == 0) {        // 16-byte vector reductions
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(w + i), "f"(v[i]), "f"(v[i + 1]),
                                     "f"(v[i + 2]), "f"(v[i + 3]) : "memory");
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) if (n + i < P.N) atomicAdd(w + i, v[i]);
                }
            }
        }
        __threadfence();
        named_bar_sync(1, 128);
        unsigned int* ctr = P.tile_counters + blockIdx.y * gridDim.x + blockIdx.x;
        if (tid == 0) hdr->pad = atomicAdd(ctr, 1u);
        named_bar_sync(1, 128);
        finish = hdr->pad == static_cast<uint32_t>(P.k_split - 1);
        if (finish) {
            __threadfence();
            if (tid == 0) *ctr = 0u;
        }
    } else {
        mbar_wait(smem_u32(&hdr->acc_full), 0, 0x53);
        tc_fence_after_sync();
    }
    for (int cg = 0; finish && cg < kBN; cg += 16) {
        float v[16];
        const int n = n0 + cg;
        if (P.k_split > 1) {
            if (!mvalid || n >= P.N) continue;
            float* w = P.ws + static_cast<long long>(m) * P.N + n;
            if (n + 16 <= P.N && (P.N & 3) == 0) {
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 t = __ldcg(reinterpret_cast<const float4*>(w + i));
                    __stcg(reinterpret_cast<float4*>(w + i), make_float4(0.f, 0.f, 0.f, 0.f));
                    v[i] = t.x; v[i + 1] = t.y; v[i + 2] = t.z; v[i + 3] = t.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    v[i] = 0.f;
                    if (n + i < P.N) { v[i] = __ldcg(w + i); __stcg(w + i, 0.f); }
                }
            }
        } else {
            tmem_ld16(tmem_base + ((warp * 32u) << 16) + cg, v);
            if (!mvalid || n >= P.N) continue;
        }
        const long long o = static_cast<long long>(m) * P.ldo + n;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            float x = v[i];
            if (n + i < P.N) {
                if (P.bias) x += P.bias[n + i];
                if (P.relu) x = fmaxf(x, 0.f);
                if (P.dropout_keep < 1.f) {      // inverted dropout, counter-based mask (seed, step, element)
                    const uint32_t h = hash32(static_cast<uint32_t>(o + i) ^ hash32(P.seed + static_cast<uint32_t>(step) * 0x9E3779B9u));
                    x = (h >> 8) * (1.f / 16777216.f) < P.dropout_keep ? x / P.dropout_keep : 0.f;
                }
            }
            v[i] = x;
        }
        if (P.out_fp32) {
            float* op = static_cast<float*>(P.out) + o;
#pragma unroll
            for (int i = 0; i < 16; ++i) if (n + i < P.N) op[i] = v[i];
        } else {
            dtf_bf16* op = static_cast<dtf_bf16*>(P.out) + o;
            if (n + 16 <= P.N && ((o & 7) == 0)) {
                reinterpret_cast<uint4*>(op)[0] = pack8(v);
                reinterpret_cast<uint4*>(op)[1] = pack8(v + 8);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) if (n + i < P.N) op[i] = __float2bfloat16(v[i]);
            }
        }
    }
}

__global__ void __launch_bounds__(kGThreads, 1)
gemm_kernel(const __grid_constant__ DtfGemm P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    GemmHeader* hdr = reinterpret_cast<GemmHeader*>(smem);
    uint8_t* stages = smem + 1024;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
    // split-K: blockIdx.z owns the K blocks [kb_begin, kb_end); a dense layer with a small batch has very few
    // output tiles (M = batch = 128 -> 8 tiles for the 3136 -> 1024 layer) but a long K loop
    const int total_kb = (P.K + kBK - 1) / kBK;
    const int kb_per = (total_kb + P.k_split - 1) / P.k_split;
    const int kb_begin = blockIdx.z * kb_per;
    const int num_kb = max(0, min(total_kb, kb_begin + kb_per) - kb_begin);
    if (tid == 0) {
        for (int i = 0; i < kGStages; ++i) { mbar_init(smem_u32(&hdr->full[i]), kGGather); mbar_init(smem_u32(&hdr->empty[i]), 1); }
        mbar_init(smem_u32(&hdr->acc_full), 1);
        mbar_fence_init();
    }
    if (warp == 4) tmem_alloc(smem_u32(&hdr->tmem_base), 128);
    griddep_launch_dependents();
    griddep_wait();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    if (warp == 4) {
        const uint32_t idesc = umma_idesc_bf16(kBM, kBN, P.a_mn_major != 0, P.b_mn_major != 0);
        for (int kb = 0; kb < num_kb; ++kb) {
            const uint32_t st = kb % kGStages;
            mbar_wait(smem_u32(&hdr->full[st]), (kb / kGStages) & 1, 0x51);
            tc_fence_after_sync();
            const uint32_t a_base = smem_u32(stages) + st * 2 * kPanelBytes, b_base = a_base + kPanelBytes;
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < kBK / 16; ++ks) {
                    const uint64_t ad = P.a_mn_major ? umma_desc(a_base + ks * 256, 128, kCsMN)
                                                     : umma_desc(a_base + ks * 2 * kCsK, kCsK, 128);
                    const uint64_t bd = P.b_mn_major ? umma_desc(b_base + ks * 256, 128, kCsMN)
                                                     : umma_desc(b_base + ks * 2 * kCsK, kCsK, 128);
                    umma_bf16(tmem_base, ad, bd, idesc, (kb | ks) != 0);
                }
                umma_commit(smem_u32(&hdr->empty[st]));
                if (kb == num_kb - 1) umma_commit(smem_u32(&hdr->acc_full));
            }
            __syncwarp();
        }
    } else if (warp > 4) {
        const int g = tid - 5 * 32;
        for (int kb = 0; kb < num_kb; ++kb) {
            const uint32_t st = kb % kGStages;
            mbar_wait(smem_u32(&hdr->empty[st]), ((kb / kGStages) & 1) ^ 1, 0x52);
            const uint32_t a_dst = smem_u32(stages) + st * 2 * kPanelBytes, b_dst = a_dst + kPanelBytes;
            load_panel(P.a, P.lda, P.a_mn_major, m0, P.M, (kb_begin + kb) * kBK, P.K, a_dst, g);
            load_panel(P.b, P.ldb, P.b_mn_major, n0, P.N, (kb_begin + kb) * kBK, P.K, b_dst, g);
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->full[st]));
        }
    } else {
        gemm_epilogue(P, hdr, tmem_base, tid, warp, lane, m0, n0, num_kb);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4) { tc_fence_after_sync(); tmem_dealloc(tmem_base, 128); }
}

// ---------------------------------------------------------------------------------------
// TMA-fed variant for two K-major row-major operands (the dense layers' forward GEMMs): no thread touches the
// operands.  One elected thread issues two cp.async.bulk.tensor.2d loads per K block (A and B panels of 128 rows
// x 64 bf16 = 128 bytes per row) through tensor maps with the 128-byte swizzle; the tensor core reads the same
// swizzled tiles (SWIZZLE_128B descriptors: 8-row atoms 1024 bytes apart, +32 bytes per K step of 16).  Rows past
// M / N and columns past K are zero-filled by the TMA unit.  Roles: 4 epilogue warps, 1 MMA warp, 1 producer warp.
// ---------------------------------------------------------------------------------------
constexpr int kTThreads = 6 * 32;
constexpr int kTStages = 4;
constexpr int kTPanel = kBM * kBK * 2;        // 16 KB, 1024-byte aligned (swizzle atom)

DTF_DEVINL void tma_load_2d(uint32_t dst_smem, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst_smem), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// K-major operand tile in the 128-byte swizzle: start address, SBO = 1024 (eight 128-byte rows), layout 2
DTF_DEVINL uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu);
    d |= static_cast<uint64_t>(1u) << 16;                          // LBO: unused for swizzled K-major layouts
    d |= static_cast<uint64_t>((1024u >> 4) & 0x3FFFu) << 32;
    d |= 1ull << 46;
    d |= 2ull << 61;
    return d;
}

__global__ void __launch_bounds__(kTThreads, 1)
gemm_tma_kernel(const __grid_constant__ DtfGemm P, const __grid_constant__ CUtensorMap map_a,
                const __grid_constant__ CUtensorMap map_b) {
    extern __shared__ __align__(1024) uint8_t smem[];
    GemmHeader* hdr = reinterpret_cast<GemmHeader*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + 256);      // kTStages full + kTStages empty barriers
    uint64_t* empty = full + kTStages;
    uint8_t* stages = smem + 1024;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
    const int total_kb = (P.K + kBK - 1) / kBK;
    const int kb_per = (total_kb + P.k_split - 1) / P.k_split;
    const int kb_begin = blockIdx.z * kb_per;
    const int num_kb = max(0, min(total_kb, kb_begin + kb_per) - kb_begin);
    if (tid == 0) {
        for (int i = 0; i < kTStages; ++i) { mbar_init(smem_u32(&full[i]), 1); mbar_init(smem_u32(&empty[i]), 1); }
        mbar_init(smem_u32(&hdr->acc_full), 1);
        mbar_fence_init();
    }
    if (warp == 5 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 4) tmem_alloc(smem_u32(&hdr->tmem_base), 128);
    griddep_launch_dependents();
    griddep_wait();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    if (warp == 5) {
        // ================= TMA producer =================
        for (int kb = 0; kb < num_kb; ++kb) {
            const uint32_t st = kb % kTStages;
            mbar_wait(smem_u32(&empty[st]), ((kb / kTStages) & 1) ^ 1, 0x54);
            if (elect_one()) {
                const uint32_t a_dst = smem_u32(stages) + st * 2 * kTPanel, b_dst = a_dst + kTPanel;
                mbar_arrive_expect_tx(smem_u32(&full[st]), 2 * kTPanel);
                tma_load_2d(a_dst, &map_a, (kb_begin + kb) * kBK, m0, smem_u32(&full[st]));
                tma_load_2d(b_dst, &map_b, (kb_begin + kb) * kBK, n0, smem_u32(&full[st]));
            }
            __syncwarp();
        }
    } else if (warp == 4) {
        // ================= MMA issuer =================
        const uint32_t idesc = umma_idesc_bf16(kBM, kBN, false, false);
        for (int kb = 0; kb < num_kb; ++kb) {
            const uint32_t st = kb % kTStages;
            mbar_wait(smem_u32(&full[st]), (kb / kTStages) & 1, 0x55);
            tc_fence_after_sync();
            const uint32_t a_base = smem_u32(stages) + st * 2 * kTPanel, b_base = a_base + kTPanel;
            if (elect_one()) {
                const uint64_t ad = umma_desc_sw128(a_base), bd = umma_desc_sw128(b_base);
#pragma unroll
                for (int ks = 0; ks < kBK / 16; ++ks)          // 16 bf16 = 32 bytes further inside the swizzle atom
                    umma_bf16(tmem_base, ad + ks * 2, bd + ks * 2, idesc, (kb | ks) != 0);
                umma_commit(smem_u32(&empty[st]));
                if (kb == num_kb - 1) umma_commit(smem_u32(&hdr->acc_full));
            }
            __syncwarp();
        }
    } else {
        gemm_epilogue(P, hdr, tmem_base, tid, warp, lane, m0, n0, num_kb);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4) { tc_fence_after_sync(); tmem_dealloc(tmem_base, 128); }
}

// ---------------------------------------------------------------------------------------
// 2x2 / stride-2 max pool, NHWC bf16 (mnist_model.py:77,87)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const dtf_bf16* __restrict__ x, dtf_bf16* __restrict__ y, int NB, int H, int W, int C) {
    griddep_launch_dependents();
    griddep_wait();
    const int Ho = H / 2, Wo = W / 2, cv = C / 8;
    const long long total = static_cast<long long>(NB) * Ho * Wo * cv;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % cv);
        long long p = i / cv;
        const int ox = static_cast<int>(p % Wo); p /= Wo;
        const int oy = static_cast<int>(p % Ho);
        const int n = static_cast<int>(p / Ho);
        float m[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) m[k] = -3.0e38f;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                float f[8];
                unpack8(ldg16(x + ((static_cast<long long>(n) * H + 2 * oy + dy) * W + 2 * ox + dx) * C + c8 * 8), f);
#pragma unroll
                for (int k = 0; k < 8; ++k) m[k] = fmaxf(m[k], f[k]);
            }
        *reinterpret_cast<uint4*>(y + ((static_cast<long long>(n) * Ho + oy) * Wo + ox) * C + c8 * 8) = pack8(m);
    }
}

// gx = gy routed to the first maximum of each window, times the ReLU mask of x (x is a post-ReLU
// activation, so mask = x > 0); also accumulates the bias gradient db[c] += sum gx.
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const dtf_bf16* __restrict__ x, const dtf_bf16* __restrict__ y, const dtf_bf16* __restrict__ gy,
                   dtf_bf16* __restrict__ gx, float* __restrict__ db, int NB, int H, int W, int C) {
    griddep_launch_dependents();
    griddep_wait();
    extern __shared__ float s_db[];
    for (int c = threadIdx.x; c < C; c += blockDim.x) s_db[c] = 0.f;
    __syncthreads();
    const int Ho = H / 2, Wo = W / 2, cv = C / 8;
    const long long total = static_cast<long long>(NB) * Ho * Wo * cv;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = static_cast<int>(i % cv);
        long long p = i / cv;
        const int ox = static_cast<int>(p % Wo); p /= Wo;
        const int oy = static_cast<int>(p % Ho);
        const int n = static_cast<int>(p / Ho);
        const long long ob = ((static_cast<long long>(n) * Ho + oy) * Wo + ox) * C + c8 * 8;
        float ym[8], g[8];
        unpack8(ldg16(y + ob), ym);
        unpack8(ldg16(gy + ob), g);
        bool taken[8] = {false, false, false, false, false, false, false, false};
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const long long ib = ((static_cast<long long>(n) * H + 2 * oy + dy) * W + 2 * ox + dx) * C + c8 * 8;
                float f[8], o[8];
                unpack8(ldg16(x + ib), f);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const bool hit = !taken[k] && f[k] == ym[k];
                    taken[k] = taken[k] || hit;
                    o[k] = (hit && f[k] > 0.f) ? g[k] : 0.f;
                    acc[k] += o[k];
                }
                *reinterpret_cast<uint4*>(gx + ib) = pack8(o);
            }
        if (db != nullptr) {
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(&s_db[c8 * 8 + k], acc[k]);
        }
    }
    __syncthreads();
    if (db != nullptr)
        for (int c = threadIdx.x; c < C; c += blockDim.x) atomicAdd(db + c, s_db[c]);
}

// g_out = g_in * (act > 0) * scale ;  db[c] += column sums      (dense + ReLU + dropout backward)
__global__ void __launch_bounds__(256)
act_mask_kernel(const dtf_bf16* __restrict__ g_in, const dtf_bf16* __restrict__ act, dtf_bf16* __restrict__ g_out,
                float* __restrict__ db, float scale, long long rows, int C) {
    griddep_launch_dependents();
    griddep_wait();
    const int cv = C / 8;
    const int c8 = blockIdx.x * blockDim.x + threadIdx.x;          // one thread = one 8-channel column strip
    if (c8 >= cv) return;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const long long o = r * C + c8 * 8;
        float g[8], a[8];
        unpack8(ldg16(g_in + o), g);
        unpack8(ldg16(act + o), a);
#pragma unroll
        for (int k = 0; k < 8; ++k) { g[k] = a[k] > 0.f ? g[k] * scale : 0.f; acc[k] += g[k]; }
        *reinterpret_cast<uint4*>(g_out + o) = pack8(g);
    }
    if (db != nullptr) {
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(db + c8 * 8 + k, acc[k]);
    }
}

}  // namespace dtf

#define DTF_CHECK_LAUNCH() (cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA)

// Tensor map of a row-major K-major bf16 matrix [rows][K] (row pitch ld elements): box 64 x 128 elements, 128-byte
// swizzle, zero fill outside.  The map only ever covers this GPU's memory (the caller's operand).
static bool make_kmajor_map(CUtensorMap* map, const void* base, int rows, int K, long long ld) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    static int tried = 0;
    if (!tried) {
        tried = 1;
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            encode = reinterpret_cast<EncodeFn>(fn);
    }
    if (encode == nullptr) return false;
    if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld * 2) & 15)) return false;
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(dtf::kBK), static_cast<cuuint32_t>(dtf::kBM)};
    const cuuint32_t estr[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

extern "C" int dtf_gemm(const DtfGemm* pin, cudaStream_t stream) {
    using namespace dtf;
    DtfGemm q = *pin;
    DtfGemm* p = &q;
    if (p->M < 1 || p->N < 1 || p->K < 1) return DTF_ERR_SHAPE;
    if ((p->lda & 7) || (p->ldb & 7) || ((uintptr_t)p->a & 15) || ((uintptr_t)p->b & 15)) return DTF_ERR_ARG;
    if (!p->a_mn_major && (p->K & 7)) return DTF_ERR_SHAPE;          // 16-byte chunks along K
    if (!p->b_mn_major && (p->K & 7)) return DTF_ERR_SHAPE;
    if (p->a_mn_major && (p->M & 7)) return DTF_ERR_SHAPE;
    if (p->b_mn_major && (p->N & 7)) return DTF_ERR_SHAPE;
    const int smem_bytes = 1024 + kGStages * 2 * kPanelBytes;
    static int set = 0;
    if (!set) {
        if (cudaFuncSetAttribute(gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes) != cudaSuccess) return DTF_ERR_CUDA;
        set = 1;
    }
    // split-K when the caller provided a (zeroed, self-cleaning) workspace and the output tiles alone cannot
    // fill the GPU: aim for ~32 CTAs (measured optimum of 1 / 32 / 64 / 120 for the MNIST dense layer; more splits
    // pay more in reductions than they save), at least two K blocks per CTA
    const int gx = (p->M + kBM - 1) / kBM, gy = (p->N + kBN - 1) / kBN;
    const int total_kb = (p->K + kBK - 1) / kBK;
    int ks = 1;
    if (p->ws != nullptr && p->tile_counters != nullptr && gx * gy < 64 && gx * gy <= p->max_tiles) {
        static int target = 0;
        if (target == 0) { const char* e = getenv("DTF_GEMM_SPLIT_TARGET"); target = e ? atoi(e) : 32; if (target < 1) target = 1; }
        ks = target / (gx * gy);
        if (ks > total_kb / 2) ks = total_kb / 2;
        if (ks > 32) ks = 32;
        if (ks < 1) ks = 1;
        const int per = (total_kb + ks - 1) / ks;
        ks = (total_kb + per - 1) / per;              // no empty splits
    }
    p->k_split = ks;
    dim3 grid(gx, gy, ks);
    // TMA-fed kernel when both operands are K-major with 16-byte aligned rows (the forward dense layers);
    // DTF_GEMM_TMA=0 keeps the thread-gather kernel for A/B measurements
    static int tma_on = -1;
    if (tma_on < 0) { const char* e = getenv("DTF_GEMM_TMA"); tma_on = e ? atoi(e) : 1; }
    if (tma_on && !p->a_mn_major && !p->b_mn_major) {
        CUtensorMap ma, mb;
        if (make_kmajor_map(&ma, p->a, p->M, p->K, p->lda) && make_kmajor_map(&mb, p->b, p->N, p->K, p->ldb)) {
            const int smem_tma = 1024 + kTStages * 2 * kTPanel;
            static int set_tma = 0;
            if (!set_tma) {
                if (cudaFuncSetAttribute(gemm_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_tma) != cudaSuccess)
                    return DTF_ERR_CUDA;
                set_tma = 1;
            }
            dtf_launch(gemm_tma_kernel, grid, dim3(kTThreads), smem_tma, stream, *p, ma, mb);
            return DTF_CHECK_LAUNCH();
        }
    }
    dtf_launch(gemm_kernel, grid, dim3(kGThreads), smem_bytes, stream, *p);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_maxpool_fwd(const dtf_bf16* x, dtf_bf16* y, int NB, int H, int W, int C, cudaStream_t stream) {
    if ((C & 7) || (H & 1) || (W & 1)) return DTF_ERR_SHAPE;
    const long long total = static_cast<long long>(NB) * (H / 2) * (W / 2) * (C / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    dtf_launch(dtf::maxpool_fwd_kernel, dim3(static_cast<int>(blocks < 1 ? 1 : blocks)), dim3(256), 0, stream, x, y, NB, H, W, C);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_maxpool_bwd(const dtf_bf16* x, const dtf_bf16* y, const dtf_bf16* gy, dtf_bf16* gx, float* db,
                               int NB, int H, int W, int C, cudaStream_t stream) {
    if ((C & 7) || (H & 1) || (W & 1) || C > 1024) return DTF_ERR_SHAPE;
    const long long total = static_cast<long long>(NB) * (H / 2) * (W / 2) * (C / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    dtf_launch(dtf::maxpool_bwd_kernel, dim3(static_cast<int>(blocks < 1 ? 1 : blocks)), dim3(256), C * sizeof(float), stream, x, y, gy, gx, db, NB, H, W, C);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_act_mask(const dtf_bf16* g_in, const dtf_bf16* act, dtf_bf16* g_out, float* db, float scale,
                            long long rows, int C, cudaStream_t stream) {
    if (C & 7) return DTF_ERR_SHAPE;
    const int cv = C / 8;
    dim3 grid((cv + 127) / 128, static_cast<unsigned>(rows < 64 ? (rows < 1 ? 1 : rows) : 64));
    dtf_launch(dtf::act_mask_kernel, grid, dim3(128), 0, stream, g_in, act, g_out, db, scale, rows, C);
    return DTF_CHECK_LAUNCH();
}
