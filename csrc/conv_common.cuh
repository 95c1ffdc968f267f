// Pieces shared by the convolution kernels (generic gather path and 3x3 halo path):
// shared-memory bookkeeping, BatchNorm constant tables, operand gather, the epilogue.
#pragma once
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

// The gather of one 16-byte chunk (8 consecutive channels) of an im2col row is split in two so that a
// thread can put the loads of several chunks in flight before it consumes the first one.
struct GatherLoad {
    uint4 raw, x;     // source chunk; second source (BN backward) when the mode needs it
    int cb;           // first channel of the chunk
    bool ok;          // false: padding / out-of-range chunk, stays zero
};

template <bool TWO>
DTF_DEVINL void gather_issue(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                             const RowCoord& rc, int jg, GatherLoad& L) {
    L.raw = make_uint4(0, 0, 0, 0);
    L.x = L.raw;
    L.ok = false;
    L.cb = 0;
    if (!rc.valid || jg >= g.kchunks_real) return;
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
        if ((ty % g.stride) != 0 || (tx % g.stride) != 0 || ty < 0 || tx < 0) return;
        sy = ty / g.stride;
        sx = tx / g.stride;
    }
    if (sy < 0 || sy >= g.Hs || sx < 0 || sx >= g.Ws) return;
    const size_t off = ((static_cast<size_t>(rc.n) * g.Hs + sy) * g.Ws + sx) * g.C + c8 * 8;
    L.raw = ldg16(g.src + off);
    if (TWO) L.x = ldg16(g.src2 + off);
    L.cb = c8 * 8;
    L.ok = true;
}

DTF_DEVINL uint4 gather_finish(const DtfGather& g, const float* p0, const float* p1, const float* p2,
                               const GatherLoad& L) {
    if (!L.ok || g.mode == DTF_PRO_NONE) return L.raw;
    float f[8];
    unpack8(L.raw, f);
    const int cb = L.cb;
    if (g.mode == DTF_PRO_BN_RELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaxf(fmaf(f[i], p0[cb + i], p1[cb + i]), 0.f);
    } else {
        float x[8];
        unpack8(L.x, x);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], p0[cb + i], fmaf(x[i], p1[cb + i], p2[cb + i]));
    }
    return pack8(f);
}

DTF_DEVINL uint4 gather16(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                          const float* p0, const float* p1, const float* p2,
                          const RowCoord& rc, int jg) {
    GatherLoad L;
    if (g.mode == DTF_PRO_BN_BWD) gather_issue<true>(g, tap_dy, tap_dx, rc, jg, L);
    else gather_issue<false>(g, tap_dy, tap_dx, rc, jg, L);
    return gather_finish(g, p0, p1, p2, L);
}

// Chunks jbeg, jbeg + 2, ... < jend of one row (global chunk index = jglobal0 + j), written to
// dst + j * kChunkBytes; BATCH loads per source are in flight per thread.
template <int BATCH, bool TWO>
DTF_DEVINL void gather_row_chunks_t(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                                    const float* p0, const float* p1, const float* p2, const RowCoord& rc,
                                    int jbeg, int jend, int jglobal0, uint32_t dst) {
    for (int j0 = jbeg; j0 < jend; j0 += 2 * BATCH) {
        GatherLoad L[BATCH];
#pragma unroll
        for (int u = 0; u < BATCH; ++u) {
            const int j = j0 + 2 * u;     // past-the-end chunks resolve to "not ok" inside gather_issue
            gather_issue<TWO>(g, tap_dy, tap_dx, rc, j < jend ? jglobal0 + j : 0x7fffffff, L[u]);
        }
#pragma unroll
        for (int u = 0; u < BATCH; ++u) {
            const int j = j0 + 2 * u;
            const uint4 r = gather_finish(g, p0, p1, p2, L[u]);
            if (j < jend) sts16(dst + j * (kTileM * 16), r);
        }
    }
}
DTF_DEVINL void gather_row_chunks(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                                  const float* p0, const float* p1, const float* p2, const RowCoord& rc,
                                  int jbeg, int jend, int jglobal0, uint32_t dst) {
    if (g.mode == DTF_PRO_BN_BWD)
        gather_row_chunks_t<4, true>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, jglobal0, dst);
    else
        gather_row_chunks_t<8, false>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, jglobal0, dst);
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


// ---------------------------------------------------------------------------------------
// epilogue of one 16-column group of one accumulator row (shared by both conv kernels)
//   v[16]  : fp32 accumulators of output pixel `obase / N`, columns cg..cg+15
//   valid  : this TMEM lane maps to a real output pixel
// ---------------------------------------------------------------------------------------
// warp butterfly of 16 per-lane partial sums per statistic -> shared bins (one add per channel per warp)
DTF_DEVINL void flush_stats16(SmemConsts* cst, float* s1, float* s2, int cg, int lane) {
    const float a1 = butterfly16(s1, lane);
    const float a2 = butterfly16(s2, lane);
    if ((lane & 1) == 0) {
        atomicAdd(&cst->bins[0][cg + butterfly_channel(lane)], a1);
        atomicAdd(&cst->bins[1][cg + butterfly_channel(lane)], a2);
    }
}

// The auxiliary tensor of the epilogue (residual, or the pre-BN input for the ReLU mask) does not
// depend on the MMA: callers prefetch it before waiting for the accumulator.
DTF_DEVINL const dtf_bf16* epilogue_aux_ptr(const DtfConvGemm& P) {
    return P.epi_mode == DTF_EPI_RESIDUAL ? P.epi_res : (P.epi_mode == DTF_EPI_RELU_MASK ? P.epi_x : nullptr);
}

// acc1/acc2 (optional): per-thread running sums of the two statistics for THIS 16-column group;
// when given, the warp butterfly is skipped here and done once per CTA by the caller.
DTF_DEVINL void epilogue_group(const DtfConvGemm& P, SmemConsts* cst, float* v, bool valid, size_t obase,
                               int cg, int lane, const uint4* aux = nullptr, float* acc1 = nullptr,
                               float* acc2 = nullptr) {
    float s2[16];
    if (P.epi_mode == DTF_EPI_RESIDUAL) {
        if (valid) {
            float r[16];
            unpack8(aux ? aux[0] : ldg16(P.epi_res + obase + cg), r);
            unpack8(aux ? aux[1] : ldg16(P.epi_res + obase + cg + 8), r + 8);
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
            unpack8(aux ? aux[0] : ldg16(P.epi_x + obase + cg), x);
            unpack8(aux ? aux[1] : ldg16(P.epi_x + obase + cg + 8), x + 8);
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
            if (acc1 != nullptr) {
#pragma unroll
                for (int i = 0; i < 16; ++i) { acc1[i] += s1[i]; acc2[i] += s2[i]; }
            } else {
                flush_stats16(cst, s1, s2, cg, lane);
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
        if (acc1 != nullptr) {
#pragma unroll
            for (int i = 0; i < 16; ++i) { acc1[i] += s1[i]; acc2[i] += s2[i]; }
        } else {
            flush_stats16(cst, s1, s2, cg, lane);
        }
    }
    if (valid) {
        uint4* o = reinterpret_cast<uint4*>(P.out + obase + cg);
        o[0] = pack8(v);
        o[1] = pack8(v + 8);
    }
}

// 32-bit division by a runtime constant (host-precomputed): q = (mulhi(n, m) + n) >> s for n < 2^31
struct FastDiv {
    uint32_t m, s, d;
};
DTF_DEVINL uint32_t fdiv(uint32_t n, const FastDiv& f) { return (__umulhi(n, f.m) + n) >> f.s; }

}  // namespace dtf
