// Implicit-GEMM convolution kernels for sm_100a: tcgen05.mma with TMEM accumulators.
//
// Replaces K1 conv fwd/dgrad/wgrad, K2 fixed_padding, K3/K4 BN-apply + ReLU
// (folded into the operand gather as a *prologue*), K5 residual add and the BN statistics
// (folded into the *epilogue*).  Reference call sites: resnet/resnet_model.py:45-92,127-320.
//
// One CTA = 13 warps:
//   warps 0-3   epilogue  : tcgen05.ld the 128 x N fp32 accumulator (one TMEM lane quadrant per
//                           warp), residual / bias / ReLU / ReLU-mask, per-channel statistics
//                           (warp butterfly -> shared bins -> one global atomic per channel per
//                           CTA), bf16 store;
//   warp  4     MMA issuer: allocates TMEM, bulk-copies (TMA, UBLKCP) the pre-packed weight
//                           matrix into shared memory once, then one elected lane issues
//                           tcgen05.mma.cta_group::1.kind::f16 per 16-wide K step and commits
//                           to mbarriers;
//   warps 5-12  gather    : build the im2col operand tile in shared memory in the canonical
//                           no-swizzle UMMA layout [k/8][128 rows][8], applying the prologue
//                           (BN scale/shift + ReLU, or the BN-backward affine combination of two
//                           tensors) on the way, zero-filling padding taps.
// Pipelines: 2 operand buffers (full/empty mbarriers) and 2 TMEM accumulators
// (acc_full/acc_empty), persistent over 128-row output tiles.
//
// The wgrad kernel reuses the same gathered tiles viewed MN-major (K = pixels): both UMMA
// operands are transposed by descriptor, D[co][tap*Cin+ci] accumulates over all tiles of the
// CTA in TMEM and is flushed with fp32 atomics into the gradient arena.
#include "common.cuh"
#include "dtf_b200.h"

namespace dtf {

constexpr int kEpiWarps = 4;
constexpr int kGatherWarps = 8;
constexpr int kThreads = (kEpiWarps + 1 + kGatherWarps) * 32;   // 416
constexpr int kGatherThreads = kGatherWarps * 32;               // 256
constexpr int kMmaWarp = kEpiWarps;
constexpr int kTileM = 128;
constexpr int kChunkBytes = kTileM * 16;                         // one [128 rows][8 bf16] chunk
constexpr int kMaxC = 256;                                       // channel bound for constants

struct __align__(16) SmemHeader {
    uint64_t full[2];
    uint64_t empty[2];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint64_t bfull;
    uint32_t tmem_base;
    uint32_t pad;
    int8_t tap_dy[32];
    int8_t tap_dx[32];
};

// Per-channel constant tables held in shared memory.
struct SmemConsts {
    float p0[kMaxC], p1[kMaxC], p2[kMaxC];        // prologue A-side
    float q0[kMaxC], q1[kMaxC], q2[kMaxC];        // prologue second operand (wgrad gy side)
    float e_scale[kMaxC], e_shift[kMaxC], e_mean[kMaxC], e_invstd[kMaxC];   // epilogue
    float bins[2][kMaxC];
};

// ---- BN helper math -------------------------------------------------------------------
// forward: y = x*scale + shift ; training statistics come as (sum, sumsq) accumulators
DTF_DEVINL void bn_fwd_consts(const DtfBnRef& bn, int c, float& scale, float& shift, float& mean, float& invstd) {
    float var;
    if (bn.use_running) {
        mean = bn.run_mean[c];
        var = bn.run_var[c];
    } else {
        mean = bn.stats[c] * bn.inv_count;
        var = fmaxf(bn.stats[bn.channels + c] * bn.inv_count - mean * mean, 0.f);
    }
    invstd = rsqrtf(var + bn.eps);
    scale = bn.gamma[c] * invstd;
    shift = bn.beta[c] - mean * scale;
}

// Fill prologue tables for one gather source.  mode 1/2: relu(x*p0+p1).  mode 3 (BN backward
// applied on the fly): g = p0*t + p1*x + p2 with
//   p0 = gamma*invstd, p1 = -gamma*invstd^2*mean(t*xhat), p2 = gamma*invstd*(mean*invstd*mean(t*xhat) - mean(t))
DTF_DEVINL void fill_prologue(const DtfGather& g, float* p0, float* p1, float* p2, int tid, int nthreads) {
    if (g.mode == DTF_PRO_NONE) return;
    for (int c = tid; c < g.C; c += nthreads) {
        float scale, shift, mean, invstd;
        bn_fwd_consts(g.bn, c, scale, shift, mean, invstd);
        if (g.mode == DTF_PRO_BN_RELU) {
            p0[c] = scale; p1[c] = shift; p2[c] = 0.f;
        } else {   // DTF_PRO_BN_BWD
            const float mt = g.bwd_stats[c] * g.bn.inv_count;                 // mean(t)      (= dbeta / M)
            const float mtx = g.bwd_stats[g.C + c] * g.bn.inv_count;          // mean(t*xhat) (= dgamma / M)
            p0[c] = scale;
            p1[c] = -scale * invstd * mtx;
            p2[c] = scale * (mean * invstd * mtx - mt);
        }
    }
}

// Moving-average update (training): decay*old + (1-decay)*batch, unbiased variance.
DTF_DEVINL void update_running(const DtfBnRef& bn, int tid, int nthreads) {
    const float m = 1.f / bn.inv_count;
    const float unbias = m > 1.f ? m / (m - 1.f) : 1.f;
    for (int c = tid; c < bn.channels; c += nthreads) {
        const float mean = bn.stats[c] * bn.inv_count;
        const float var = fmaxf(bn.stats[bn.channels + c] * bn.inv_count - mean * mean, 0.f);
        bn.run_mean[c] = bn.decay * bn.run_mean[c] + (1.f - bn.decay) * mean;
        bn.run_var[c] = bn.decay * bn.run_var[c] + (1.f - bn.decay) * var * unbias;
    }
}

// ---- operand gather -------------------------------------------------------------------
struct RowCoord { int n, oy, ox; bool valid; };

DTF_DEVINL RowCoord row_coord(int m, int M, int Ho, int Wo) {
    RowCoord r;
    r.valid = m < M;
    const int mm = r.valid ? m : 0;
    const int hw = Ho * Wo;
    r.n = mm / hw;
    const int rem = mm - r.n * hw;
    r.oy = rem / Wo;
    r.ox = rem - r.oy * Wo;
    return r;
}

// Returns the 16-byte chunk (8 consecutive channels) of the im2col row for global chunk jg.
DTF_DEVINL uint4 gather16(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                          const float* p0, const float* p1, const float* p2,
                          const RowCoord& rc, int jg) {
    const uint4 zero = make_uint4(0, 0, 0, 0);
    if (!rc.valid || jg >= g.kchunks_real) return zero;
    const int tap = jg >> g.cpt_shift;
    const int c8 = jg & ((1 << g.cpt_shift) - 1);
    int sy, sx;
    if (!g.transposed) {
        sy = rc.oy * g.stride + tap_dy[tap];
        sx = rc.ox * g.stride + tap_dx[tap];
    } else {
        // gradient gather of a strided conv: source = (o + pad - r) / stride when divisible
        const int ty = rc.oy - tap_dy[tap];
        const int tx = rc.ox - tap_dx[tap];
        if ((ty % g.stride) != 0 || (tx % g.stride) != 0 || ty < 0 || tx < 0) return zero;
        sy = ty / g.stride;
        sx = tx / g.stride;
    }
    if (sy < 0 || sy >= g.Hs || sx < 0 || sx >= g.Ws) return zero;
    const size_t off = ((static_cast<size_t>(rc.n) * g.Hs + sy) * g.Ws + sx) * g.C + c8 * 8;
    const uint4 raw = ldg16(g.src + off);
    if (g.mode == DTF_PRO_NONE) return raw;
    float f[8];
    unpack8(raw, f);
    const int cb = c8 * 8;
    if (g.mode == DTF_PRO_BN_RELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaxf(fmaf(f[i], p0[cb + i], p1[cb + i]), 0.f);
    } else {
        float x[8];
        unpack8(ldg16(g.src2 + off), x);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], p0[cb + i], fmaf(x[i], p1[cb + i], p2[cb + i]));
    }
    return pack8(f);
}

DTF_DEVINL void fill_taps(SmemHeader* h, const DtfGather& g, int tid) {
    if (tid < g.KH * g.KW && tid < 32) {
        const int r = tid / g.KW, s = tid - r * g.KW;
        // forward: source = o*stride + r - pad ; transposed: t = o + pad - r  => o - (r - pad)
        h->tap_dy[tid] = static_cast<int8_t>(r - g.pad);
        h->tap_dx[tid] = static_cast<int8_t>(s - g.pad);
    }
    if (tid == 31) { h->tap_dy[31] = 0; h->tap_dx[31] = 0; }   // identity tap (1x1 gathers)
}

// 16-value butterfly: on return lanes hold in v[0] the 32-lane sum of channel
// ((lane>>4)&1)*8 + ((lane>>3)&1)*4 + ((lane>>2)&1)*2 + ((lane>>1)&1).
DTF_DEVINL float butterfly16(float* v, int lane) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool hi = lane & 16;
        const float send = hi ? v[i] : v[i + 8];
        const float keep = hi ? v[i + 8] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool hi = lane & 8;
        const float send = hi ? v[i] : v[i + 4];
        const float keep = hi ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const bool hi = lane & 4;
        const float send = hi ? v[i] : v[i + 2];
        const float keep = hi ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    {
        const bool hi = lane & 2;
        const float send = hi ? v[0] : v[1];
        const float keep = hi ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    return v[0];
}
DTF_DEVINL int butterfly_channel(int lane) {
    return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}

// =========================================================================================
// forward / dgrad kernel:  out[M][N] = epilogue( gather(A)[M][K] * Wpacked[N][K]^T )
// =========================================================================================
__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ DtfConvGemm P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    SmemHeader* hdr = reinterpret_cast<SmemHeader*>(smem);
    SmemConsts* cst = reinterpret_cast<SmemConsts*>(smem + 1024);
    uint8_t* smem_b = smem + P.smem_b_off;
    uint8_t* smem_a = smem + P.smem_a_off;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int N = P.N;
    const int M = P.NB * P.Ho * P.Wo;
    const int num_tiles = (M + kTileM - 1) / kTileM;
    const int kb_chunks = P.kb_chunks;
    const int num_kb = (P.kchunks + kb_chunks - 1) / kb_chunks;
    const uint32_t acc_cols = tmem_cols_pow2(2 * N);

    // ---- one-time setup ----
    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&hdr->full[i]), kGatherThreads);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
            mbar_init(smem_u32(&hdr->acc_full[i]), 1);
            mbar_init(smem_u32(&hdr->acc_empty[i]), kEpiWarps * 32);
        }
        mbar_init(smem_u32(&hdr->bfull), 1);
        mbar_fence_init();
    }
    fill_taps(hdr, P.a, tid);
    fill_prologue(P.a, cst->p0, cst->p1, cst->p2, tid, kThreads);
    if (P.epi_mode == DTF_EPI_RELU_MASK) {
        for (int c = tid; c < N; c += kThreads)
            bn_fwd_consts(P.epi_bn, c, cst->e_scale[c], cst->e_shift[c], cst->e_mean[c], cst->e_invstd[c]);
    }
    for (int c = tid; c < 2 * kMaxC; c += kThreads) (&cst->bins[0][0])[c] = 0.f;
    if (warp == kMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    if (blockIdx.x == 0 && P.a.update_running && P.a.mode == DTF_PRO_BN_RELU && !P.a.bn.use_running)
        update_running(P.a.bn, tid, kThreads);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;

    if (warp == kMmaWarp) {
        // ================= MMA issuer =================
        if (lane == 0) {
            const uint32_t wbytes = static_cast<uint32_t>(P.kchunks) * N * 16u;
            mbar_arrive_expect_tx(smem_u32(&hdr->bfull), wbytes);
            bulk_g2s(smem_u32(smem_b), P.wpk, wbytes, smem_u32(&hdr->bfull));
            mbar_wait(smem_u32(&hdr->bfull), 0, 1);
            const uint32_t idesc = umma_idesc_bf16(kTileM, N, false, false);
            uint32_t it = 0, local_tile = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
                const uint32_t a = local_tile & 1;
                mbar_wait(smem_u32(&hdr->acc_empty[a]), ((local_tile >> 1) & 1) ^ 1, 2);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + a * N;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const uint32_t buf = it & 1;
                    mbar_wait(smem_u32(&hdr->full[buf]), (it >> 1) & 1, 3);
                    tc_fence_after_sync();
                    const int nch = min(kb_chunks, P.kchunks - kb * kb_chunks);
                    const uint32_t a_base = smem_u32(smem_a) + buf * kb_chunks * kChunkBytes;
                    const uint32_t b_base = smem_u32(smem_b) + static_cast<uint32_t>(kb * kb_chunks) * N * 16u;
                    for (int ks = 0; ks < nch / 2; ++ks) {
                        const uint64_t adesc = umma_desc(a_base + ks * 2 * kChunkBytes, kChunkBytes, 128);
                        const uint64_t bdesc = umma_desc(b_base + ks * 2 * N * 16, N * 16, 128);
                        umma_bf16(d_tmem, adesc, bdesc, idesc, (kb | ks) != 0);
                    }
                    umma_commit(smem_u32(&hdr->empty[buf]));
                    if (kb == num_kb - 1) umma_commit(smem_u32(&hdr->acc_full[a]));
                }
            }
        }
        __syncwarp();
    } else if (warp > kMmaWarp) {
        // ================= gather producers =================
        const int g = tid - (kMmaWarp + 1) * 32;
        const int row = g & (kTileM - 1);
        const int half = g >> 7;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const RowCoord rc = row_coord(tile * kTileM + row, M, P.Ho, P.Wo);
            for (int kb = 0; kb < num_kb; ++kb, ++it) {
                const uint32_t buf = it & 1;
                mbar_wait(smem_u32(&hdr->empty[buf]), ((it >> 1) & 1) ^ 1, 4);
                const int nch = min(kb_chunks, P.kchunks - kb * kb_chunks);
                const uint32_t dst = smem_u32(smem_a) + buf * kb_chunks * kChunkBytes + row * 16;
                int j = half;
                for (; j + 6 < nch; j += 8) {   // 4 chunks in flight per thread
                    uint4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        v[u] = gather16(P.a, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc,
                                        kb * kb_chunks + j + 2 * u);
#pragma unroll
                    for (int u = 0; u < 4; ++u) sts16(dst + (j + 2 * u) * kChunkBytes, v[u]);
                }
                for (; j < nch; j += 2)
                    sts16(dst + j * kChunkBytes,
                          gather16(P.a, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, kb * kb_chunks + j));
                fence_proxy_async_smem();
                mbar_arrive(smem_u32(&hdr->full[buf]));
            }
        }
    } else {
        // ================= epilogue =================
        uint32_t local_tile = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            const uint32_t a = local_tile & 1;
            mbar_wait(smem_u32(&hdr->acc_full[a]), (local_tile >> 1) & 1, 5);
            tc_fence_after_sync();
            const int m = tile * kTileM + warp * 32 + lane;
            const bool valid = m < M;
            const size_t obase = static_cast<size_t>(valid ? m : 0) * N;
            for (int cg = 0; cg < N; cg += 16) {
                float v[16];
                tmem_ld16(tmem_base + ((warp * 32u) << 16) + a * N + cg, v);
                float s2[16];
                if (P.epi_mode == DTF_EPI_RESIDUAL) {
                    if (valid) {
                        float r[16];
                        unpack8(ldg16(P.epi_res + obase + cg), r);
                        unpack8(ldg16(P.epi_res + obase + cg + 8), r + 8);
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] += r[i];
                    }
                } else if (P.epi_mode == DTF_EPI_BIAS_RELU || P.epi_mode == DTF_EPI_BIAS) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        v[i] += P.epi_bias[cg + i];
                        if (P.epi_mode == DTF_EPI_BIAS_RELU) v[i] = fmaxf(v[i], 0.f);
                    }
                } else if (P.epi_mode == DTF_EPI_RELU_MASK) {
                    float x[16];
                    if (valid) {
                        unpack8(ldg16(P.epi_x + obase + cg), x);
                        unpack8(ldg16(P.epi_x + obase + cg + 8), x + 8);
                    }
                    float old[16];
                    if (P.epi_accumulate && valid) {
                        unpack8(ldg16(P.out + obase + cg), old);
                        unpack8(ldg16(P.out + obase + cg + 8), old + 8);
                    }
                    float s1[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float xv = valid ? x[i] : 0.f;
                        const bool on = valid && fmaf(xv, cst->e_scale[cg + i], cst->e_shift[cg + i]) > 0.f;
                        // statistics only see THIS launch's contribution; an accumulating launch adds
                        // its own share to the same global accumulators (the sums are linear).
                        const float t = on ? v[i] : 0.f;
                        s1[i] = t;
                        s2[i] = t * (xv - cst->e_mean[cg + i]) * cst->e_invstd[cg + i];
                        v[i] = (P.epi_accumulate && valid) ? t + old[i] : t;
                    }
                    if (P.epi_stats_out) {
                        const float a1 = butterfly16(s1, lane);
                        const float a2 = butterfly16(s2, lane);
                        if ((lane & 1) == 0) {
                            atomicAdd(&cst->bins[0][cg + butterfly_channel(lane)], a1);
                            atomicAdd(&cst->bins[1][cg + butterfly_channel(lane)], a2);
                        }
                    }
                }
                if (P.epi_mode != DTF_EPI_RELU_MASK && P.epi_stats_out) {
                    float s1[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        s1[i] = valid ? v[i] : 0.f;
                        s2[i] = s1[i] * s1[i];
                    }
                    const float a1 = butterfly16(s1, lane);
                    const float a2 = butterfly16(s2, lane);
                    if ((lane & 1) == 0) {
                        atomicAdd(&cst->bins[0][cg + butterfly_channel(lane)], a1);
                        atomicAdd(&cst->bins[1][cg + butterfly_channel(lane)], a2);
                    }
                }
                if (valid) {
                    uint4* o = reinterpret_cast<uint4*>(P.out + obase + cg);
                    o[0] = pack8(v);
                    o[1] = pack8(v + 8);
                }
            }
            tc_fence_before_sync();
            mbar_arrive(smem_u32(&hdr->acc_empty[a]));
        }
        if (P.epi_stats_out) {
            named_bar_sync(1, kEpiWarps * 32);
            for (int c = tid; c < N; c += kEpiWarps * 32) {
                atomicAdd(P.epi_stats_out + c, cst->bins[0][c]);
                atomicAdd(P.epi_stats_out + N + c, cst->bins[1][c]);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, acc_cols);
    }
}

// =========================================================================================
// wgrad kernel:  dW[co][j*8+e] += sum over pixels  gy'[pixel][co] * im2col(act(x))[pixel][j*8+e]
//   blockIdx.y selects a contiguous range of im2col chunks (<= 32 chunks = 256 columns).
// =========================================================================================
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ DtfConvWgrad P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    SmemHeader* hdr = reinterpret_cast<SmemHeader*>(smem);
    SmemConsts* cst = reinterpret_cast<SmemConsts*>(smem + 1024);
    uint8_t* smem_ab = smem + P.smem_a_off;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int M = P.NB * P.Ho * P.Wo;
    const int num_tiles = (M + kTileM - 1) / kTileM;
    const int j0 = blockIdx.y * P.chunks_per_cta;
    const int nj = min(P.chunks_per_cta, P.kchunks - j0);            // real+pad chunks of this CTA (even)
    const int ncols = nj * 8;                                         // UMMA N
    const int gy_chunks = P.Cout / 8;
    const uint32_t buf_bytes = static_cast<uint32_t>(gy_chunks + P.chunks_per_cta) * kChunkBytes;
    const uint32_t acc_cols = tmem_cols_pow2(ncols);

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&hdr->full[i]), kGatherThreads);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
        }
        mbar_init(smem_u32(&hdr->acc_full[0]), 1);
        mbar_fence_init();
    }
    fill_taps(hdr, P.x, tid);
    fill_prologue(P.x, cst->p0, cst->p1, cst->p2, tid, kThreads);
    fill_prologue(P.gy, cst->q0, cst->q1, cst->q2, tid, kThreads);
    if (warp == kMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    const bool has_work = blockIdx.x < num_tiles;

    if (warp == kMmaWarp) {
        if (lane == 0 && has_work) {
            // A = gy' viewed MN-major (M index = channel, 8 per 16-byte row), B = im2col MN-major.
            const uint32_t idesc = umma_idesc_bf16(kTileM, ncols, true, true);
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const uint32_t buf = it & 1;
                mbar_wait(smem_u32(&hdr->full[buf]), (it >> 1) & 1, 6);
                tc_fence_after_sync();
                const uint32_t a_base = smem_u32(smem_ab) + buf * buf_bytes;
                const uint32_t b_base = a_base + gy_chunks * kChunkBytes;
                for (int ks = 0; ks < kTileM / 16; ++ks) {
                    // K step = 16 pixels = 2 groups of 8 rows (LBO = 128); channel chunks SBO apart
                    const uint64_t adesc = umma_desc(a_base + ks * 256, 128, kChunkBytes);
                    const uint64_t bdesc = umma_desc(b_base + ks * 256, 128, kChunkBytes);
                    umma_bf16(tmem_base, adesc, bdesc, idesc, (it | ks) != 0);
                }
                umma_commit(smem_u32(&hdr->empty[buf]));
            }
            umma_commit(smem_u32(&hdr->acc_full[0]));
        }
        __syncwarp();
    } else if (warp > kMmaWarp) {
        const int g = tid - (kMmaWarp + 1) * 32;
        const int row = g & (kTileM - 1);
        const int half = g >> 7;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const uint32_t buf = it & 1;
            const RowCoord rc = row_coord(tile * kTileM + row, M, P.Ho, P.Wo);
            mbar_wait(smem_u32(&hdr->empty[buf]), ((it >> 1) & 1) ^ 1, 7);
            const uint32_t a_dst = smem_u32(smem_ab) + buf * buf_bytes + row * 16;
            const uint32_t b_dst = a_dst + gy_chunks * kChunkBytes;
            // gy' chunks: a 1x1 "gather" at the output pixel itself
            for (int j = half; j < gy_chunks; j += 2)
                sts16(a_dst + j * kChunkBytes,
                      gather16(P.gy, hdr->tap_dy + 31, hdr->tap_dx + 31, cst->q0, cst->q1, cst->q2, rc, j));
            int j = half;
            for (; j + 6 < nj; j += 8) {
                uint4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    v[u] = gather16(P.x, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, j0 + j + 2 * u);
#pragma unroll
                for (int u = 0; u < 4; ++u) sts16(b_dst + (j + 2 * u) * kChunkBytes, v[u]);
            }
            for (; j < nj; j += 2)
                sts16(b_dst + j * kChunkBytes,
                      gather16(P.x, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, j0 + j));
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->full[buf]));
        }
    } else if (has_work) {
        mbar_wait(smem_u32(&hdr->acc_full[0]), 0, 8);
        tc_fence_after_sync();
        const int co = warp * 32 + lane;                   // TMEM lane = output channel
        const int kreal = P.kchunks_real * 8;
        for (int cg = 0; cg < ncols; cg += 8) {
            float v[8];
            tmem_ld8(tmem_base + ((warp * 32u) << 16) + cg, v);   // whole warp executes the load
            if (co < P.Cout) {
                const int k = j0 * 8 + cg;
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (k + i < kreal) atomicAdd(P.dw + static_cast<size_t>(co) * kreal + k + i, v[i]);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, acc_cols);
    }
}

}  // namespace dtf

// =========================================================================================
// host launchers (C ABI)
// =========================================================================================
static int g_num_sms = 0;
static int num_sms() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

static int ilog2_exact(int v) {
    int s = 0;
    while ((1 << s) < v) ++s;
    return ((1 << s) == v) ? s : -1;
}

static int prep_gather(DtfGather* g) {
    if (g->C % 8 != 0) return DTF_ERR_SHAPE;
    const int cpt = g->C / 8;
    g->cpt_shift = ilog2_exact(cpt);
    if (g->cpt_shift < 0) return DTF_ERR_SHAPE;
    if (g->KH * g->KW > 31) return DTF_ERR_SHAPE;
    if (g->C > dtf::kMaxC) return DTF_ERR_SHAPE;
    g->kchunks_real = g->KH * g->KW * cpt;
    return DTF_OK;
}

extern "C" int dtf_conv_gemm(DtfConvGemm* p, cudaStream_t stream) {
    using namespace dtf;
    int rc = prep_gather(&p->a);
    if (rc != DTF_OK) return rc;
    if (p->N % 16 != 0 || p->N < 16 || p->N > 256) return DTF_ERR_SHAPE;
    p->kchunks = (p->a.kchunks_real + 1) & ~1;
    // operand buffer: as many chunks as fit next to the weights, even, <= kchunks
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    const int b_off = (header + 1023) & ~1023;
    const int b_bytes = p->kchunks * p->N * 16;
    const int a_off = (b_off + b_bytes + 1023) & ~1023;
    const int budget = 227 * 1024 - a_off;
    int kb = (budget / (2 * kChunkBytes)) & ~1;
    if (kb > p->kchunks) kb = p->kchunks;
    if (kb > 24) kb = 24;                      // smaller blocks -> finer-grained overlap
    if (kb < 2) return DTF_ERR_SMEM;
    p->kb_chunks = kb;
    p->smem_b_off = b_off;
    p->smem_a_off = a_off;
    const int smem_bytes = a_off + 2 * kb * kChunkBytes;
    const int M = p->NB * p->Ho * p->Wo;
    const int tiles = (M + kTileM - 1) / kTileM;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    static int max_set = 0;
    if (smem_bytes > max_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return DTF_ERR_CUDA;
        max_set = 227 * 1024;
    }
    conv_gemm_kernel<<<grid, kThreads, smem_bytes, stream>>>(*p);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

extern "C" int dtf_conv_wgrad(DtfConvWgrad* p, cudaStream_t stream) {
    using namespace dtf;
    int rc = prep_gather(&p->x);
    if (rc != DTF_OK) return rc;
    // gy side: 1x1 gather at the output pixel
    p->gy.KH = p->gy.KW = 1; p->gy.stride = 1; p->gy.pad = 0; p->gy.transposed = 0;
    p->gy.Hs = p->Ho; p->gy.Ws = p->Wo; p->gy.C = p->Cout;
    rc = prep_gather(&p->gy);
    if (rc != DTF_OK) return rc;
    if (p->Cout % 8 != 0 || p->Cout > 128) return DTF_ERR_SHAPE;
    p->kchunks_real = p->x.kchunks_real;
    p->kchunks = (p->kchunks_real + 1) & ~1;
    // columns per CTA: <= 24 chunks (192 columns), even, N % 16 == 0 holds for even chunk counts
    int per = p->kchunks;
    int ny = 1;
    while (per > 24) { ++ny; per = (((p->kchunks + ny - 1) / ny) + 1) & ~1; }
    p->chunks_per_cta = per;
    ny = (p->kchunks + per - 1) / per;
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    const int a_off = (header + 1023) & ~1023;
    const int buf_bytes = (p->Cout / 8 + per) * kChunkBytes;
    // the A descriptor always spans 16 channel chunks: keep 32 KB of slack after the last buffer
    const int smem_bytes = a_off + 2 * buf_bytes + 16 * kChunkBytes;
    if (smem_bytes > 227 * 1024) return DTF_ERR_SMEM;
    p->smem_a_off = a_off;
    const int M = p->NB * p->Ho * p->Wo;
    const int tiles = (M + kTileM - 1) / kTileM;
    int gx = num_sms() / ny;
    if (gx < 1) gx = 1;
    if (gx > tiles) gx = tiles;
    static int set = 0;
    if (!set) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return DTF_ERR_CUDA;
        set = 1;
    }
    dim3 grid(gx, ny);
    conv_wgrad_kernel<<<grid, kThreads, smem_bytes, stream>>>(*p);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
