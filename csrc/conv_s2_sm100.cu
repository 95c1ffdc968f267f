// Input gradient of a stride-2 convolution (3x3 pad 1, or the 1x1 projection shortcut), split by
// output-pixel parity so that no multiply ever sees a structural zero.
//
// For a stride-2 conv the gradient of input pixel (iy, ix) only receives the filter taps whose row
// index r satisfies (iy + pad - r) even (same for columns).  With a 3x3 / pad-1 filter:
//     iy even: r = 1            (source row iy/2)
//     iy odd : r = 0 or r = 2   (source rows (iy+1)/2 and (iy-1)/2)
// so the four parity classes (iy&1, ix&1) use 1, 2, 2 and 4 taps -- 9 in total instead of the
// 4 x 9 a dense transposed gather walks through (three quarters of which are zero chunks).  A tile
// here is 128 pixels of ONE class; its K extent is (taps of the class) x Cout.  The 1x1 projection
// only has class (0, 0): the other three quarters of the gradient are untouched, which is exactly
// right for the accumulating launch that adds the shortcut gradient on top of the main path.
//
// Reference math: the gradient TF computes for `conv2d_fixed_padding(strides=2)`
// (resnet/resnet_model.py:76-92, projection shortcut 127-135).
//
// Warp roles, pipelines, shared-memory layout and the epilogue are those of conv_sm100.cu; the
// weights stay in the dgrad packing [tap][Cout/8][Cin][8] and the B descriptor simply starts at the
// tap it needs.  All gather loads of a tile are issued before the first one is consumed.
#include <cstdio>
#include <cstdlib>
#include "conv_common.cuh"

namespace dtf {

struct __align__(16) S2Header {
    uint64_t full[2];
    uint64_t empty[2];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint64_t bfull;
    uint32_t tmem_base;
    uint32_t pad;
    int8_t ntaps[4];
    int8_t dy[16], dx[16], widx[16];   // [class * 4 + tap]
};

struct S2Row { int n, a, b; bool valid; };

DTF_DEVINL S2Row s2_row(int m, int Mc, int Hh, int Wh) {
    const RowCoord c = row_coord(m, Mc, Hh, Wh);
    S2Row r;
    r.n = c.n; r.a = c.oy; r.b = c.ox; r.valid = c.valid;
    return r;
}

// TPB taps of one row starting at tap t0: every chunk of every tap is loaded (both sources when TWO) before the
// first one goes through the prologue and into shared memory.
template <int L, bool TWO, int TPB>
DTF_DEVINL void s2_batch_t(const DtfGather& G, const S2Header* hdr, const SmemConsts* cst, int cls, const S2Row& rc,
                           int rowbase, uint32_t dst, int t0) {
    constexpr int PER = 1 << L;
    uint4 v[TPB * PER], x[TPB * PER];
    uint32_t okmask = 0;
#pragma unroll
    for (int u = 0; u < TPB; ++u) {
        const int ti = cls * 4 + t0 + u;
        const int sy = rc.a + hdr->dy[ti], sx = rc.b + hdr->dx[ti];
        const bool ok = rc.valid && sy < G.Hs && sx < G.Ws;
        const int off = (rowbase + sy * G.Ws + sx) * G.C;
#pragma unroll
        for (int c = 0; c < PER; ++c) {
            const int c8 = c;
            v[u * PER + c] = ok ? ldg16(G.src + off + c8 * 8) : make_uint4(0, 0, 0, 0);
            if (TWO) x[u * PER + c] = ok ? ldg16(G.src2 + off + c8 * 8) : make_uint4(0, 0, 0, 0);
        }
        okmask |= (ok ? 1u : 0u) << u;
    }
#pragma unroll
    for (int u = 0; u < TPB; ++u) {
#pragma unroll
        for (int c = 0; c < PER; ++c) {
            const int c8 = c;
            uint4 r = v[u * PER + c];
            if (TWO && ((okmask >> u) & 1)) {
                const int cb = c8 * 8;
                float f[8], xx[8];
                unpack8(r, f);
                unpack8(x[u * PER + c], xx);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    f[i] = fmaf(f[i], cst->p0[cb + i], fmaf(xx[i], cst->p1[cb + i], cst->p2[cb + i]));
                r = pack8(f);
            }
            sts16(dst + (((t0 + u) << L) + c8) * kChunkBytes, r);
        }
    }
}

// One row of the operand tile, all taps of its parity class (1, 2 or 4).  The loads of a batch are issued before
// the first is consumed; batches are sized to the class so that no dead tap costs instructions.
template <int L, bool TWO>
DTF_DEVINL void s2_gather_t(const DtfGather& G, const S2Header* hdr, const SmemConsts* cst, int cls, const S2Row& rc,
                            int rowbase, uint32_t dst) {
    constexpr int PER = 1 << L;
    constexpr int BUDGET = TWO ? 4 : 8;
    const int nt = hdr->ntaps[cls];
    if (nt == 1) {
        s2_batch_t<L, TWO, 1>(G, hdr, cst, cls, rc, rowbase, dst, 0);
    } else if (nt == 2) {
        if (2 * PER <= BUDGET) {
            s2_batch_t<L, TWO, (2 * PER <= BUDGET ? 2 : 1)>(G, hdr, cst, cls, rc, rowbase, dst, 0);
        } else {
            s2_batch_t<L, TWO, 1>(G, hdr, cst, cls, rc, rowbase, dst, 0);
            s2_batch_t<L, TWO, 1>(G, hdr, cst, cls, rc, rowbase, dst, 1);
        }
    } else {
        if (4 * PER <= BUDGET) {
            s2_batch_t<L, TWO, (4 * PER <= BUDGET ? 4 : 1)>(G, hdr, cst, cls, rc, rowbase, dst, 0);
        } else if (2 * PER <= BUDGET) {
            s2_batch_t<L, TWO, (2 * PER <= BUDGET ? 2 : 1)>(G, hdr, cst, cls, rc, rowbase, dst, 0);
            s2_batch_t<L, TWO, (2 * PER <= BUDGET ? 2 : 1)>(G, hdr, cst, cls, rc, rowbase, dst, 2);
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) s2_batch_t<L, TWO, 1>(G, hdr, cst, cls, rc, rowbase, dst, t);
        }
    }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_dgrad_s2_kernel(const __grid_constant__ DtfConvGemm P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    S2Header* hdr = reinterpret_cast<S2Header*>(smem);
    SmemConsts* cst = reinterpret_cast<SmemConsts*>(smem + 1024);
    uint8_t* smem_b = smem + P.smem_b_off;
    uint8_t* smem_a = smem + P.smem_a_off;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int N = P.N;
    const DtfGather& G = P.a;
    const int Hh = P.Ho >> 1, Wh = P.Wo >> 1;
    const int Mc = P.NB * Hh * Wh;                              // pixels per parity class
    const int tiles_per_class = (Mc + kTileM - 1) / kTileM;
    const int nclass = G.KH == 3 ? 4 : 1;
    const int num_tiles = nclass * tiles_per_class;
    const int cpt = 1 << G.cpt_shift;                           // 8-channel chunks per tap
    const int cap = P.kb_chunks;                                // chunk capacity of one operand buffer
    const uint32_t acc_cols = tmem_cols_pow2(2 * N);

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&hdr->full[i]), kGatherGroup);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
            mbar_init(smem_u32(&hdr->acc_full[i]), 1);
            mbar_init(smem_u32(&hdr->acc_empty[i]), kEpiWarps * 32);
        }
        mbar_init(smem_u32(&hdr->bfull), 1);
        mbar_fence_init();
        if (G.KH == 3) {
            for (int py = 0; py < 2; ++py)
                for (int px = 0; px < 2; ++px) {
                    const int c = py * 2 + px;
                    int n = 0;
                    for (int r = 0; r < 3; ++r) {
                        if (((r + 1) & 1) != py) continue;      // r == 1 for even rows, r in {0, 2} for odd
                        for (int s = 0; s < 3; ++s) {
                            if (((s + 1) & 1) != px) continue;
                            hdr->dy[c * 4 + n] = static_cast<int8_t>(r == 0 ? 1 : 0);
                            hdr->dx[c * 4 + n] = static_cast<int8_t>(s == 0 ? 1 : 0);
                            hdr->widx[c * 4 + n] = static_cast<int8_t>(r * 3 + s);
                            ++n;
                        }
                    }
                    hdr->ntaps[c] = static_cast<int8_t>(n);
                }
        } else {
            hdr->ntaps[0] = 1;
            hdr->dy[0] = hdr->dx[0] = hdr->widx[0] = 0;
        }
    }
    for (int c = tid; c < 2 * kMaxC; c += kThreads) (&cst->bins[0][0])[c] = 0.f;
    if (warp == kMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    griddep_launch_dependents();
    griddep_wait();
    fill_prologue(G, cst->p0, cst->p1, cst->p2, tid, kThreads);
    if (P.epi_mode == DTF_EPI_RELU_MASK) {
        for (int c = tid; c < N; c += kThreads)
            bn_fwd_consts(P.epi_bn, c, cst->e_scale[c], cst->e_shift[c], cst->e_mean[c], cst->e_invstd[c]);
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;

    if (warp == kMmaWarp) {
        // ================= MMA issuer =================
        const uint32_t wbytes = static_cast<uint32_t>(P.kchunks) * N * 16u;
        if (elect_one()) {
            mbar_arrive_expect_tx(smem_u32(&hdr->bfull), wbytes);
            bulk_g2s(smem_u32(smem_b), P.wpk, wbytes, smem_u32(&hdr->bfull));
        }
        __syncwarp();
        mbar_wait(smem_u32(&hdr->bfull), 0, 21);
        const uint32_t idesc = umma_idesc_bf16(kTileM, N, false, false);
        const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int cls = tile / tiles_per_class;
            const int nt = hdr->ntaps[cls];
            const uint32_t buf = it & 1;
            mbar_wait(smem_u32(&hdr->acc_empty[buf]), ((it >> 1) & 1) ^ 1, 22);
            mbar_wait(smem_u32(&hdr->full[buf]), (it >> 1) & 1, 23);
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem_base + buf * N;
            const uint32_t a_base = a0 + buf * cap * kChunkBytes;
            for (int t = 0; t < nt; ++t) {
                const uint32_t b_tap = b0 + static_cast<uint32_t>(hdr->widx[cls * 4 + t]) * cpt * N * 16u;
                for (int k2 = 0; k2 < cpt; k2 += 2) {
                    const uint64_t adesc = umma_desc(a_base + (t * cpt + k2) * kChunkBytes, kChunkBytes, 128);
                    const uint64_t bdesc = umma_desc(b_tap + k2 * N * 16, N * 16, 128);
                    if (elect_one()) umma_bf16(d_tmem, adesc, bdesc, idesc, (t | k2) != 0);
                }
            }
            if (elect_one()) {
                umma_commit(smem_u32(&hdr->empty[buf]));
                umma_commit(smem_u32(&hdr->acc_full[buf]));
            }
            __syncwarp();
        }
    } else if (warp > kMmaWarp) {
        // ================= gather producers =================
        const int g = tid - (kMmaWarp + 1) * 32;
        const int row = g & (kTileM - 1);
        const uint32_t group = g >> 7;                       // two groups work on alternate tiles
        const bool two = G.mode == DTF_PRO_BN_BWD;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            if ((it & 1) != group) continue;
            const int cls = tile / tiles_per_class;
            const S2Row rc = s2_row((tile - cls * tiles_per_class) * kTileM + row, Mc, Hh, Wh);
            mbar_wait(smem_u32(&hdr->empty[group]), ((it >> 1) & 1) ^ 1, 24);
            const uint32_t dst = smem_u32(smem_a) + group * cap * kChunkBytes + row * 16;
            const int rowbase = rc.n * G.Hs * G.Ws;
#define DTF_S2_CASE(LL)                                                                                   \
    case LL:                                                                                              \
        if (two) s2_gather_t<LL, true>(G, hdr, cst, cls, rc, rowbase, dst);                         \
        else s2_gather_t<LL, false>(G, hdr, cst, cls, rc, rowbase, dst);                            \
        break;
            switch (G.cpt_shift) {
                DTF_S2_CASE(1)
                DTF_S2_CASE(2)
                DTF_S2_CASE(3)
            }
#undef DTF_S2_CASE
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->full[group]));
        }
    } else {
        // ================= epilogue =================
        const dtf_bf16* auxp = epilogue_aux_ptr(P);
        const bool reg_stats = P.epi_stats_out != nullptr && N <= 16;   // see conv_sm100.cu
        float acc1[16], acc2[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { acc1[i] = 0.f; acc2[i] = 0.f; }
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int cls = tile / tiles_per_class;
            const S2Row rc = s2_row((tile - cls * tiles_per_class) * kTileM + warp * 32 + lane, Mc, Hh, Wh);
            const int oy = 2 * rc.a + (cls >> 1), ox = 2 * rc.b + (cls & 1);
            const size_t obase = ((static_cast<size_t>(rc.n) * P.Ho + oy) * P.Wo + ox) * N;
            uint4 aux[8];                                          // N <= 64
#pragma unroll
            for (int i = 0; i < 8; ++i)
                aux[i] = (auxp != nullptr && rc.valid && i * 8 < N) ? ldg16(auxp + obase + i * 8)
                                                                    : make_uint4(0, 0, 0, 0);
            const uint32_t a = it & 1;
            mbar_wait(smem_u32(&hdr->acc_full[a]), (it >> 1) & 1, 25);
            tc_fence_after_sync();
            if (reg_stats) {
                float v[16];
                tmem_ld16(tmem_base + ((warp * 32u) << 16) + a * N, v);
                epilogue_group(P, cst, v, rc.valid, obase, 0, lane, aux, acc1, acc2);
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int cg = q * 16;
                    if (cg < N) {
                        float v[16];
                        tmem_ld16(tmem_base + ((warp * 32u) << 16) + a * N + cg, v);
                        epilogue_group(P, cst, v, rc.valid, obase, cg, lane, aux + 2 * q);
                    }
                }
            }
            tc_fence_before_sync();
            mbar_arrive(smem_u32(&hdr->acc_empty[a]));
        }
        if (reg_stats) flush_stats16(cst, acc1, acc2, 0, lane);
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

}  // namespace dtf

extern "C" int dtf_sm_count();

// Returns 1 = launched, 0 = not applicable (caller falls back to the dense transposed gather), < 0 = error.
// `p->a` has been through prep_gather and p->kchunks is set.
extern "C" int dtf_conv_s2_dgrad_try(const DtfConvGemm* p, cudaStream_t stream) {
    using namespace dtf;
    static int disabled = -1;
    if (disabled < 0) disabled = getenv("DTF_CONV_GENERIC") != nullptr ? 1 : 0;
    if (disabled) return 0;
    const DtfGather& g = p->a;
    if (g.dtype != DTF_BF16 || !g.transposed || g.stride != 2 || g.KH != g.KW) return 0;
    const bool k3 = g.KH == 3 && g.pad == 1, k1 = g.KH == 1 && g.pad == 0;
    if (!k3 && !k1) return 0;
    if (k1 && !(p->epi_mode == DTF_EPI_RELU_MASK && p->epi_accumulate)) return 0;   // 3/4 of `out` is not written
    if ((p->Ho & 1) || (p->Wo & 1) || g.Hs * 2 != p->Ho || g.Ws * 2 != p->Wo) return 0;
    if (g.cpt_shift < 1 || g.cpt_shift > 3 || p->N > 64 || g.mode == DTF_PRO_BN_RELU) return 0;
    const int cpt = 1 << g.cpt_shift;
    DtfConvGemm q = *p;
    q.kb_chunks = (k3 ? 4 : 1) * cpt;
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    q.smem_b_off = (header + 1023) & ~1023;
    q.smem_a_off = (q.smem_b_off + q.kchunks * q.N * 16 + 1023) & ~1023;
    const int smem_bytes = q.smem_a_off + 2 * q.kb_chunks * kChunkBytes;
    if (smem_bytes > 227 * 1024) return 0;
    static int set = 0;
    if (!set) {
        const cudaError_t e =
            cudaFuncSetAttribute(conv_dgrad_s2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) {
            fprintf(stderr, "[dtf] conv_dgrad_s2: cudaFuncSetAttribute: %s\n", cudaGetErrorString(e));
            return DTF_ERR_CUDA;
        }
        set = 1;
    }
    const int sms = dtf_sm_count();
    const int Mc = q.NB * (q.Ho / 2) * (q.Wo / 2);
    const int tiles = (k3 ? 4 : 1) * ((Mc + kTileM - 1) / kTileM);
    const int grid = tiles < sms ? tiles : sms;
    dtf_launch(conv_dgrad_s2_kernel, dim3(grid), dim3(kThreads), smem_bytes, stream, q);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[dtf] conv_dgrad_s2 launch (grid %d, smem %d): %s\n", grid, smem_bytes, cudaGetErrorString(e));
        return DTF_ERR_CUDA;
    }
    return 1;
}
