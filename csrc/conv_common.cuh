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
constexpr int kGatherGroup = kGatherThreads / 2;                // one thread per tile row; two groups alternate
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
    float wbins[8][2][64];     // halo kernels: private statistics bins of each epilogue warp (no atomics, N <= 64)
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
    if (((hw & (hw - 1)) | (Wo & (Wo - 1))) == 0) {      // powers of two (every CIFAR layer): shifts only
        const int shw = 31 - __clz(hw), sw = 31 - __clz(Wo);
        r.n = mm >> shw;
        const int rem = mm & (hw - 1);
        r.oy = rem >> sw;
        r.ox = rem & (Wo - 1);
        return r;
    }
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

// E = elements per 16-byte chunk: 8 (bf16) or 4 (fp32 storage, tf32 MMA).  The structs type every tensor pointer
// as dtf_bf16*; elem_ptr<E> turns an element offset into the address for the launch's element type.
template <int E>
DTF_DEVINL const dtf_bf16* elem_ptr(const dtf_bf16* base, size_t elem) {
    return reinterpret_cast<const dtf_bf16*>(reinterpret_cast<const char*>(base) + elem * (16 / E));
}
template <int E>
DTF_DEVINL dtf_bf16* elem_ptr(dtf_bf16* base, size_t elem) {
    return reinterpret_cast<dtf_bf16*>(reinterpret_cast<char*>(base) + elem * (16 / E));
}
template <int E>
DTF_DEVINL void unpack_chunk(const uint4& q, float* f) {
    if (E == 8) {
        unpack8(q, f);
    } else {
        f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y);
        f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    }
}
template <int E>
DTF_DEVINL uint4 pack_chunk(const float* f) {
    if (E == 8) return pack8(f);
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}
// 16 consecutive elements <-> 16 floats (two chunks of bf16, four of fp32)
template <int E>
DTF_DEVINL void load16(const dtf_bf16* base, size_t elem, float* f) {
#pragma unroll
    for (int i = 0; i < 16 / E; ++i) unpack_chunk<E>(ldg16(elem_ptr<E>(base, elem + i * E)), f + i * E);
}
template <int E>
DTF_DEVINL void store16(dtf_bf16* base, size_t elem, const float* f) {
    uint4* o = reinterpret_cast<uint4*>(elem_ptr<E>(base, elem));
#pragma unroll
    for (int i = 0; i < 16 / E; ++i) o[i] = pack_chunk<E>(f + i * E);
}

template <bool TWO, int E = 8>
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
    const size_t off = ((static_cast<size_t>(rc.n) * g.Hs + sy) * g.Ws + sx) * g.C + c8 * E;
    L.raw = ldg16(elem_ptr<E>(g.src, off));
    if (TWO) L.x = ldg16(elem_ptr<E>(g.src2, off));
    L.cb = c8 * E;
    L.ok = true;
}

template <int E = 8>
DTF_DEVINL uint4 gather_finish(const DtfGather& g, const float* p0, const float* p1, const float* p2,
                               const GatherLoad& L) {
    if (!L.ok || g.mode == DTF_PRO_NONE) return L.raw;
    float f[E];
    unpack_chunk<E>(L.raw, f);
    const int cb = L.cb;
    if (g.mode == DTF_PRO_BN_RELU) {
#pragma unroll
        for (int i = 0; i < E; ++i) f[i] = fmaxf(fmaf(f[i], p0[cb + i], p1[cb + i]), 0.f);
    } else {
        float x[E];
        unpack_chunk<E>(L.x, x);
#pragma unroll
        for (int i = 0; i < E; ++i) f[i] = fmaf(f[i], p0[cb + i], fmaf(x[i], p1[cb + i], p2[cb + i]));
    }
    return pack_chunk<E>(f);
}

DTF_DEVINL uint4 gather16(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                          const float* p0, const float* p1, const float* p2,
                          const RowCoord& rc, int jg) {
    GatherLoad L;
    if (g.mode == DTF_PRO_BN_BWD) gather_issue<true>(g, tap_dy, tap_dx, rc, jg, L);
    else gather_issue<false>(g, tap_dy, tap_dx, rc, jg, L);
    return gather_finish<8>(g, p0, p1, p2, L);
}

// Shared-memory address of chunk jj (counted from the first chunk of the call) of one row.
//   SW = false: chunk-major, [chunk][128 rows][16 bytes] (K-major operands, and MN-major bf16 operands);
//                `dst` = buffer + row * 16.
//   SW = true : the one MN-major layout kind::tf32 accepts (measured with bench/probe_umma.py: no-swizzle, 32B, 64B
//                and plain 128B MN-major descriptors read zeros for tf32): layout type 1, "128-byte swizzle with a
//                32-byte base".  Rows of 128 bytes = 32 channels at a 128-byte pitch, blocks of 32 channels 16 KB
//                apart, and within a row the 32-byte unit index is XOR-ed with (row & 3).
//                `dst` = buffer + row * 128, `sw` = (row & 3) << 1 in units of 16-byte chunks.
constexpr uint32_t kSwBlockBytes = kTileM * 128;       // one 32-channel block of a 128-row tile
template <bool SW>
DTF_DEVINL uint32_t chunk_addr(uint32_t dst, int jj, uint32_t sw) {
    if (SW) return dst + (static_cast<uint32_t>(jj >> 3) * kSwBlockBytes) + (((static_cast<uint32_t>(jj) & 7u) ^ sw) << 4);
    return dst + jj * (kTileM * 16);
}

// ---- row gather, tap-major ---------------------------------------------------------------
// Chunks j in [jbeg, jend) of one im2col row, written to dst + (j - jbeg) * 2 KB.  One thread owns the
// whole row of a tile; the two gather groups (4 warps each) work on alternate tiles so that two tiles'
// loads are in flight per CTA.
// The source pixel of a tap is resolved ONCE per tap (not per chunk), all address arithmetic is 32-bit,
// strides are powers of two, and the loads of a batch of taps are issued before any is consumed.
//   L = log2(chunks per tap); TWO = second source (BN backward) is read as well.
struct RowPre {
    int base;        // n * Hs * Ws
    int oy, ox;      // output pixel (already multiplied by the stride for a forward gather)
    bool valid;
};
DTF_DEVINL RowPre row_pre(const DtfGather& g, const RowCoord& rc) {
    RowPre r;
    r.base = rc.n * g.Hs * g.Ws;
    r.oy = g.transposed ? rc.oy : rc.oy * g.stride;
    r.ox = g.transposed ? rc.ox : rc.ox * g.stride;
    r.valid = rc.valid;
    return r;
}

// One batch of TPB taps starting at t0 (taps beyond t_end are dead).  Gather warps share an issue port
// with one other warp only, so the serial instruction stream per tile is what bounds small-K layers:
// the caller picks TPB to match the live tap count instead of predicating a fixed-size batch.
template <int L, bool TWO, int TPB, int E = 8, bool SW = false>
DTF_DEVINL void gather_batch_t(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                               const float* p0, const float* p1, const float* p2, const RowPre& rp,
                               int jbeg, int jend, uint32_t dst, int t0, int t_end, uint32_t sw = 0) {
    constexpr int CPT = 1 << L;
    constexpr int PER = CPT;                                  // chunks per tap
    constexpr int TSTEP = 1;
    const int smask = g.stride - 1;
    const int sshift = g.stride >> 1;                         // stride in {1, 2}
    uint4 v[TPB * PER], x[TPB * PER];
    uint32_t okmask = 0;
#pragma unroll
    for (int u = 0; u < TPB; ++u) {
        const int t = t0 + u * TSTEP;
        bool ok = rp.valid && t < t_end;
        int sy, sx;
        const int tt = t < 32 ? t : 31;
        if (!g.transposed) {
            sy = rp.oy + tap_dy[tt];
            sx = rp.ox + tap_dx[tt];
        } else {
            const int ty = rp.oy - tap_dy[tt], tx = rp.ox - tap_dx[tt];
            ok = ok && ((ty | tx) & smask) == 0;
            sy = ty >> sshift;
            sx = tx >> sshift;
        }
        ok = ok && static_cast<unsigned>(sy) < static_cast<unsigned>(g.Hs) &&
             static_cast<unsigned>(sx) < static_cast<unsigned>(g.Ws);
        const int off = (rp.base + sy * g.Ws + sx) * g.C;
#pragma unroll
        for (int c = 0; c < PER; ++c) {
            const int c8 = c;
            const int j = (t << L) + c8;
            const bool live = ok && j >= jbeg && j < jend && j < g.kchunks_real;
            v[u * PER + c] = live ? ldg16(elem_ptr<E>(g.src, off + c8 * E)) : make_uint4(0, 0, 0, 0);
            if (TWO) x[u * PER + c] = live ? ldg16(elem_ptr<E>(g.src2, off + c8 * E)) : make_uint4(0, 0, 0, 0);
            okmask |= (live ? 1u : 0u) << (u * PER + c);
        }
    }
#pragma unroll
    for (int u = 0; u < TPB; ++u) {
        const int t = t0 + u * TSTEP;
#pragma unroll
        for (int c = 0; c < PER; ++c) {
            const int c8 = c;
            const int j = (t << L) + c8;
            uint4 r = v[u * PER + c];
            if (g.mode != DTF_PRO_NONE && ((okmask >> (u * PER + c)) & 1)) {
                float f[E];
                unpack_chunk<E>(r, f);
                const int cb = c8 * E;
                if (!TWO) {
#pragma unroll
                    for (int i = 0; i < E; ++i) f[i] = fmaxf(fmaf(f[i], p0[cb + i], p1[cb + i]), 0.f);
                } else {
                    float xx[E];
                    unpack_chunk<E>(x[u * PER + c], xx);
#pragma unroll
                    for (int i = 0; i < E; ++i) f[i] = fmaf(f[i], p0[cb + i], fmaf(xx[i], p1[cb + i], p2[cb + i]));
                }
                r = pack_chunk<E>(f);
            }
            if (t < t_end && j >= jbeg && j < jend) sts16(chunk_addr<SW>(dst, j - jbeg, sw), r);
        }
    }
}

template <int L, bool TWO, int E = 8, bool SW = false>
DTF_DEVINL void gather_taps_t(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                              const float* p0, const float* p1, const float* p2, const RowPre& rp,
                              int jbeg, int jend, uint32_t dst, uint32_t sw = 0) {
    constexpr int PER = 1 << L;
    constexpr int BUDGET = TWO ? 4 : 8;                       // 16-byte loads in flight per source
    constexpr int TMAX = BUDGET / PER > 0 ? BUDGET / PER : 1; // taps per full batch
    constexpr int TSTEP = 1;
    int t = jbeg >> L;
    const int t_end = ((jend - 1) >> L) + 1;
#define DTF_BATCH(N)                                                                                       \
    {                                                                                                      \
        gather_batch_t<L, TWO, N, E, SW>(g, tap_dy, tap_dx, p0, p1, p2, rp, jbeg, jend, dst, t, t_end, sw); \
        t += N * TSTEP;                                                                                    \
    }
    while (t < t_end) {
        const int rem = (t_end - t + TSTEP - 1) / TSTEP;
        if (TMAX >= 8 && rem >= 5) DTF_BATCH(TMAX >= 8 ? 8 : 1)
        else if (TMAX >= 4 && rem >= 3) DTF_BATCH(TMAX >= 4 ? 4 : 1)
        else if (TMAX >= 2 && rem >= 2) DTF_BATCH(TMAX >= 2 ? 2 : 1)
        else DTF_BATCH(1)
    }
#undef DTF_BATCH
}

// Fallback (strides > 2): one chunk at a time.
DTF_DEVINL void gather_chunks_slow(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                                   const float* p0, const float* p1, const float* p2, const RowCoord& rc,
                                   int jbeg, int jend, uint32_t dst) {
    for (int j = jbeg; j < jend; ++j)
        sts16(dst + (j - jbeg) * (kTileM * 16), gather16(g, tap_dy, tap_dx, p0, p1, p2, rc, j));
}

// Wide channel counts (128 / 256: the 1x1 convolutions of bottleneck blocks): the chunks of a row are taken in
// batches whose loads are all in flight before the first one is consumed (8 x 16 bytes per source, 4 with the
// second source of the BN-backward prologue).
template <bool TWO, int E = 8, bool SW = false>
DTF_DEVINL void gather_chunks_wide(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                                   const float* p0, const float* p1, const float* p2, const RowCoord& rc,
                                   int jbeg, int jend, uint32_t dst, uint32_t sw = 0) {
    constexpr int B = TWO ? 4 : 8;
    for (int j = jbeg; j < jend; j += B) {
        GatherLoad L[B];
#pragma unroll
        for (int u = 0; u < B; ++u)
            gather_issue<TWO, E>(g, tap_dy, tap_dx, rc, j + u < jend ? j + u : g.kchunks_real, L[u]);
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (j + u < jend) sts16(chunk_addr<SW>(dst, j + u - jbeg, sw), gather_finish<E>(g, p0, p1, p2, L[u]));
    }
}

template <int E = 8, bool SW = false>
DTF_DEVINL void gather_row(const DtfGather& g, const int8_t* tap_dy, const int8_t* tap_dx,
                           const float* p0, const float* p1, const float* p2, const RowCoord& rc,
                           int jbeg, int jend, uint32_t dst, uint32_t sw = 0) {
    if (E != 8) {          // fp32 tensors: tap-major batches up to 8 chunks (32 channels) per tap, else chunk batches
        const bool two32 = g.mode == DTF_PRO_BN_BWD;
        if (g.cpt_shift > 3 || g.stride > 2) {
            if (two32) gather_chunks_wide<true, E, SW>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, dst, sw);
            else gather_chunks_wide<false, E, SW>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, dst, sw);
            return;
        }
        const RowPre rp32 = row_pre(g, rc);
#define DTF_GATHER_CASE32(LL)                                                                                     \
    case LL:                                                                                                      \
        if (two32) gather_taps_t<LL, true, E, SW>(g, tap_dy, tap_dx, p0, p1, p2, rp32, jbeg, jend, dst, sw);       \
        else gather_taps_t<LL, false, E, SW>(g, tap_dy, tap_dx, p0, p1, p2, rp32, jbeg, jend, dst, sw);            \
        break;
        switch (g.cpt_shift) {
            DTF_GATHER_CASE32(1)
            DTF_GATHER_CASE32(2)
            DTF_GATHER_CASE32(3)
        }
#undef DTF_GATHER_CASE32
        return;
    }
    if (g.stride > 2) {
        gather_chunks_slow(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, dst);
        return;
    }
    if (g.cpt_shift > 3) {
        if (g.mode == DTF_PRO_BN_BWD) gather_chunks_wide<true>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, dst);
        else gather_chunks_wide<false>(g, tap_dy, tap_dx, p0, p1, p2, rc, jbeg, jend, dst);
        return;
    }
    const RowPre rp = row_pre(g, rc);
    const bool two = g.mode == DTF_PRO_BN_BWD;
#define DTF_GATHER_CASE(LL)                                                                             \
    case LL:                                                                                            \
        if (two) gather_taps_t<LL, true>(g, tap_dy, tap_dx, p0, p1, p2, rp, jbeg, jend, dst);            \
        else gather_taps_t<LL, false>(g, tap_dy, tap_dx, p0, p1, p2, rp, jbeg, jend, dst);               \
        break;
    switch (g.cpt_shift) {
        DTF_GATHER_CASE(0)
        DTF_GATHER_CASE(1)
        DTF_GATHER_CASE(2)
        DTF_GATHER_CASE(3)
    }
#undef DTF_GATHER_CASE
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
// `wb` (optional): this warp's private bins [2][64] -- after the butterfly each even lane owns one channel, so a
// plain read-modify-write replaces the shared-memory atomic (eight warps hitting the same 16 addresses every tile
// made those atomics the most expensive part of the epilogue).
DTF_DEVINL void flush_stats16(SmemConsts* cst, float* s1, float* s2, int cg, int lane, float* wb = nullptr) {
    const float a1 = butterfly16(s1, lane);
    const float a2 = butterfly16(s2, lane);
    if ((lane & 1) == 0) {
        const int c = cg + butterfly_channel(lane);
        if (wb != nullptr) {
            wb[c] += a1;
            wb[64 + c] += a2;
        } else {
            atomicAdd(&cst->bins[0][c], a1);
            atomicAdd(&cst->bins[1][c], a2);
        }
    }
}

// The auxiliary tensor of the epilogue (residual, or the pre-BN input for the ReLU mask) does not
// depend on the MMA: callers prefetch it before waiting for the accumulator.
DTF_DEVINL const dtf_bf16* epilogue_aux_ptr(const DtfConvGemm& P) {
    return P.epi_mode == DTF_EPI_RESIDUAL ? P.epi_res : (P.epi_mode == DTF_EPI_RELU_MASK ? P.epi_x : nullptr);
}

// acc1/acc2 (optional): per-thread running sums of the two statistics for THIS 16-column group;
// when given, the warp butterfly is skipped here and done once per CTA by the caller.
template <int E = 8>
DTF_DEVINL void epilogue_group(const DtfConvGemm& P, SmemConsts* cst, float* v, bool valid, size_t obase,
                               int cg, int lane, const uint4* aux = nullptr, float* acc1 = nullptr,
                               float* acc2 = nullptr, float* wb = nullptr) {
    float s2[16];
    if (P.epi_mode == DTF_EPI_RESIDUAL) {
        if (valid) {
            float r[16];
            if (E == 8 && aux) { unpack8(aux[0], r); unpack8(aux[1], r + 8); }
            else load16<E>(P.epi_res, obase + cg, r);
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
            if (E == 8 && aux) { unpack8(aux[0], x); unpack8(aux[1], x + 8); }
            else load16<E>(P.epi_x, obase + cg, x);
        }
        float old[16];
        if (P.epi_accumulate && valid) load16<E>(P.out, obase + cg, old);
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
                flush_stats16(cst, s1, s2, cg, lane, wb);
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
            flush_stats16(cst, s1, s2, cg, lane, wb);
        }
    }
    if (valid) store16<E>(P.out, obase + cg, v);
}

// 32-bit division by a runtime constant (host-precomputed): q = (mulhi(n, m) + n) >> s for n < 2^31
struct FastDiv {
    uint32_t m, s, d;
};
DTF_DEVINL uint32_t fdiv(uint32_t n, const FastDiv& f) { return (__umulhi(n, f.m) + n) >> f.s; }

}  // namespace dtf
