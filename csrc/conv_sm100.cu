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
// CTA in TMEM and is flushed with fp32 atomics into the gradient arena.  More than 128 output
// channels (bottleneck blocks) are split over blockIdx.z.
//
// Both kernels are templates on E, the elements per 16-byte chunk: E = 8 for bf16 tensors
// (kind::f16) and E = 4 for fp32 tensors multiplied as tf32 (kind::tf32; the reference's default
// dtype).  K-major operands keep their byte layout (two chunks = 32 bytes of K per MMA either
// way).  MN-major tf32 operands exist in one layout only -- 128-byte rows with a 32-byte swizzle
// base, descriptor layout type 1 (bench/probe_umma.py) -- so the fp32 wgrad kernel writes its
// tiles through chunk_addr<true> (conv_common.cuh) instead of the chunk-major form.
#include "conv_common.cuh"

namespace dtf {

// =========================================================================================
// forward / dgrad kernel:  out[M][N] = epilogue( gather(A)[M][K] * Wpacked[N][K]^T )
// =========================================================================================
// E = elements per 16-byte chunk: 8 = bf16 tensors (kind::f16), 4 = fp32 tensors (kind::tf32).  Both consume two
// chunks (32 bytes of K per row) per MMA, so the shared-memory layouts and descriptors are the same.
template <int E>
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
            mbar_init(smem_u32(&hdr->full[i]), kGatherGroup);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
            mbar_init(smem_u32(&hdr->acc_full[i]), 1);
            mbar_init(smem_u32(&hdr->acc_empty[i]), kEpiWarps * 32);
        }
        mbar_init(smem_u32(&hdr->bfull), 1);
        mbar_fence_init();
    }
    fill_taps(hdr, P.a, tid);
    for (int c = tid; c < 2 * kMaxC; c += kThreads) (&cst->bins[0][0])[c] = 0.f;
    if (warp == kMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    griddep_launch_dependents();
    griddep_wait();
    fill_prologue(P.a, cst->p0, cst->p1, cst->p2, tid, kThreads);
    if (P.epi_mode == DTF_EPI_RELU_MASK) {
        for (int c = tid; c < N; c += kThreads)
            bn_fwd_consts(P.epi_bn, c, cst->e_scale[c], cst->e_shift[c], cst->e_mean[c], cst->e_invstd[c]);
    }
    if (blockIdx.x == 0 && P.a.update_running && P.a.mode == DTF_PRO_BN_RELU && !P.a.bn.use_running)
        update_running(P.a.bn, tid, kThreads);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;

    if (warp == kMmaWarp) {
        // ================= MMA issuer (warp-uniform loop, one elected lane issues) =================
        const uint32_t wbytes = static_cast<uint32_t>(P.kchunks) * N * 16u;
        if (elect_one()) {
            mbar_arrive_expect_tx(smem_u32(&hdr->bfull), wbytes);
            bulk_g2s(smem_u32(smem_b), P.wpk, wbytes, smem_u32(&hdr->bfull));
        }
        __syncwarp();
        mbar_wait(smem_u32(&hdr->bfull), 0, 1);
        const uint32_t idesc = E == 8 ? umma_idesc_bf16(kTileM, N, false, false) : umma_idesc_tf32(kTileM, N, false, false);
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
                    if (elect_one()) {
                        if (E == 8) umma_bf16(d_tmem, adesc, bdesc, idesc, (kb | ks) != 0);
                        else umma_tf32(d_tmem, adesc, bdesc, idesc, (kb | ks) != 0);
                    }
                }
                if (elect_one()) {
                    umma_commit(smem_u32(&hdr->empty[buf]));
                    if (kb == num_kb - 1) umma_commit(smem_u32(&hdr->acc_full[a]));
                }
                __syncwarp();
            }
        }
    } else if (warp > kMmaWarp) {
        // ================= gather producers =================
        const int g = tid - (kMmaWarp + 1) * 32;
        const int row = g & (kTileM - 1);
        const uint32_t group = g >> 7;                       // operand buffer this thread's group fills
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const RowCoord rc = row_coord(tile * kTileM + row, M, P.Ho, P.Wo);
            for (int kb = 0; kb < num_kb; ++kb, ++it) {
                if ((it & 1) != group) continue;
                mbar_wait(smem_u32(&hdr->empty[group]), ((it >> 1) & 1) ^ 1, 4);
                const int nch = min(kb_chunks, P.kchunks - kb * kb_chunks);
                const uint32_t dst = smem_u32(smem_a) + group * kb_chunks * kChunkBytes + row * 16;
                gather_row<E>(P.a, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, kb * kb_chunks,
                              kb * kb_chunks + nch, dst);
                fence_proxy_async_smem();
                mbar_arrive(smem_u32(&hdr->full[group]));
            }
        }
    } else {
        // ================= epilogue =================
        // N <= 16: the per-channel statistics stay in registers across the CTA's tiles (one warp
        // butterfly per CTA instead of one per tile -- the epilogue warp's instruction stream is what
        // bounds the small-K layers).
        const bool reg_stats = P.epi_stats_out != nullptr && N <= 16;
        float acc1[16], acc2[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { acc1[i] = 0.f; acc2[i] = 0.f; }
        uint32_t local_tile = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            const uint32_t a = local_tile & 1;
            mbar_wait(smem_u32(&hdr->acc_full[a]), (local_tile >> 1) & 1, 5);
            tc_fence_after_sync();
            const int m = tile * kTileM + warp * 32 + lane;
            const bool valid = m < M;
            const size_t obase = static_cast<size_t>(valid ? m : 0) * N;
            if (reg_stats) {
                float v[16];
                tmem_ld16(tmem_base + ((warp * 32u) << 16) + a * N, v);
                epilogue_group<E>(P, cst, v, valid, obase, 0, lane, nullptr, acc1, acc2);
            } else {
                for (int cg = 0; cg < N; cg += 16) {
                    float v[16];
                    tmem_ld16(tmem_base + ((warp * 32u) << 16) + a * N + cg, v);
                    epilogue_group<E>(P, cst, v, valid, obase, cg, lane);
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

// =========================================================================================
// wgrad kernel:  dW[co][j*8+e] += sum over pixels  gy'[pixel][co] * im2col(act(x))[pixel][j*8+e]
//   blockIdx.y selects a contiguous range of im2col chunks (<= 32 chunks = 256 columns).
// =========================================================================================
template <int E>
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ DtfConvWgrad P) {
    constexpr int kCoChunks = 128 / E;          // channel chunks of one 128-channel M block
    constexpr int kStepRows = E == 8 ? 16 : 8;  // pixels (K) per MMA: 32 bytes of K per channel
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
    const int ncols = nj * E;                                         // UMMA N
    // blockIdx.z selects 128 output channels (Cout up to 256: the expanding 1x1 convolutions of bottleneck blocks)
    const int co_chunk0 = blockIdx.z * kCoChunks;
    const int gy_chunks = min(P.Cout / E - co_chunk0, kCoChunks);
    // bf16: both operands chunk-major ([chunk][128 pixels][16 bytes]), gy' first.  fp32: MN-major tf32 operands exist
    // only in the swizzled 128-byte-row layout (conv_common.cuh: chunk_addr), blocks of 32 channels / columns of
    // 16 KB each, im2col blocks first, then the gy' blocks.
    const uint32_t b_blocks = (P.chunks_per_cta + 7) >> 3;
    const uint32_t a_blocks = (min(P.Cout / E, kCoChunks) + 7) >> 3;
    const uint32_t buf_bytes = E == 8 ? static_cast<uint32_t>(min(P.Cout / E, kCoChunks) + P.chunks_per_cta) * kChunkBytes
                                      : (a_blocks + b_blocks) * kSwBlockBytes;
    const uint32_t acc_cols = tmem_cols_pow2(ncols);

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&hdr->full[i]), kGatherGroup + kTileM);   // x rows (one gather group) + gy' rows (epilogue warps)
            mbar_init(smem_u32(&hdr->empty[i]), 1);
        }
        mbar_init(smem_u32(&hdr->acc_full[0]), 1);
        mbar_fence_init();
    }
    fill_taps(hdr, P.x, tid);
    if (warp == kMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    griddep_launch_dependents();
    griddep_wait();
    fill_prologue(P.x, cst->p0, cst->p1, cst->p2, tid, kThreads);
    fill_prologue(P.gy, cst->q0, cst->q1, cst->q2, tid, kThreads);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    const bool has_work = blockIdx.x < num_tiles;

    if (warp == kMmaWarp) {
        if (has_work) {
            // A = gy' viewed MN-major (M index = channel, 8 per 16-byte row), B = im2col MN-major.
            const uint32_t idesc = E == 8 ? umma_idesc_bf16(kTileM, ncols, true, true) : umma_idesc_tf32(kTileM, ncols, true, true);
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const uint32_t buf = it & 1;
                mbar_wait(smem_u32(&hdr->full[buf]), (it >> 1) & 1, 6);
                tc_fence_after_sync();
                const uint32_t buf_base = smem_u32(smem_ab) + buf * buf_bytes;
                const uint32_t a_base = E == 8 ? buf_base : buf_base + b_blocks * kSwBlockBytes;
                const uint32_t b_base = E == 8 ? buf_base + gy_chunks * kChunkBytes : buf_base;
                for (int ks = 0; ks < kTileM / kStepRows; ++ks) {
                    uint64_t adesc, bdesc;
                    if (E == 8) {
                        // K step = 16 pixels = 2 groups of 8 rows (LBO = 128); channel chunks SBO apart
                        adesc = umma_desc(a_base + ks * 256, 128, kChunkBytes);
                        bdesc = umma_desc(b_base + ks * 256, 128, kChunkBytes);
                    } else {
                        // K step = 8 pixels = 8 rows of 128 bytes (two 4-row swizzle atoms, SBO = 512 apart);
                        // 32-channel blocks LBO = 16 KB apart; layout type 1 in descriptor bits 61..63
                        adesc = umma_desc(a_base + ks * 1024, kSwBlockBytes, 512) | (1ull << 61);
                        bdesc = umma_desc(b_base + ks * 1024, kSwBlockBytes, 512) | (1ull << 61);
                    }
                    if (elect_one()) {
                        if (E == 8) umma_bf16(tmem_base, adesc, bdesc, idesc, (it | ks) != 0);
                        else umma_tf32(tmem_base, adesc, bdesc, idesc, (it | ks) != 0);
                    }
                }
                if (elect_one()) umma_commit(smem_u32(&hdr->empty[buf]));
                __syncwarp();
            }
            if (elect_one()) umma_commit(smem_u32(&hdr->acc_full[0]));
        }
        __syncwarp();
    } else if (warp > kMmaWarp) {
        const int g = tid - (kMmaWarp + 1) * 32;
        const int row = g & (kTileM - 1);
        const uint32_t group = g >> 7;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            if ((it & 1) != group) continue;
            const RowCoord rc = row_coord(tile * kTileM + row, M, P.Ho, P.Wo);
            mbar_wait(smem_u32(&hdr->empty[group]), ((it >> 1) & 1) ^ 1, 7);
            if (E == 8) {
                const uint32_t b_dst = smem_u32(smem_ab) + group * buf_bytes + gy_chunks * kChunkBytes + row * 16;
                gather_row<E>(P.x, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, j0, j0 + nj, b_dst);
            } else {
                const uint32_t b_dst = smem_u32(smem_ab) + group * buf_bytes + row * 128;
                gather_row<E, true>(P.x, hdr->tap_dy, hdr->tap_dx, cst->p0, cst->p1, cst->p2, rc, j0, j0 + nj, b_dst,
                                    (row & 3) << 1);
            }
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->full[group]));
        }
    } else if (has_work) {
        // The epilogue warps only flush at the very end: meanwhile they load the gy' rows of every tile (one row
        // per thread), in parallel with the gather groups that build the im2col operand.
        {
            const int row = tid;                               // 4 warps = 128 rows
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const uint32_t buf = it & 1;
                const RowCoord rc = row_coord(tile * kTileM + row, M, P.Ho, P.Wo);
                mbar_wait(smem_u32(&hdr->empty[buf]), ((it >> 1) & 1) ^ 1, 9);
                if (E == 8) {
                    const uint32_t a_dst = smem_u32(smem_ab) + buf * buf_bytes + row * 16;
                    gather_row<E>(P.gy, hdr->tap_dy + 31, hdr->tap_dx + 31, cst->q0, cst->q1, cst->q2, rc, co_chunk0,
                                  co_chunk0 + gy_chunks, a_dst);
                } else {
                    const uint32_t a_dst = smem_u32(smem_ab) + buf * buf_bytes + b_blocks * kSwBlockBytes + row * 128;
                    gather_row<E, true>(P.gy, hdr->tap_dy + 31, hdr->tap_dx + 31, cst->q0, cst->q1, cst->q2, rc, co_chunk0,
                                        co_chunk0 + gy_chunks, a_dst, (row & 3) << 1);
                }
                fence_proxy_async_smem();
                mbar_arrive(smem_u32(&hdr->full[buf]));
            }
        }
        mbar_wait(smem_u32(&hdr->acc_full[0]), 0, 8);
        tc_fence_after_sync();
        const int co = co_chunk0 * E + warp * 32 + lane;   // TMEM lane = output channel
        const int kreal = P.kchunks_real * E;
        for (int cg = 0; cg < ncols; cg += 8) {
            float v[8];
            tmem_ld8(tmem_base + ((warp * 32u) << 16) + cg, v);   // whole warp executes the load
            if (co < P.Cout) {
                const int k = j0 * E + cg;
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
// SMs the persistent kernels may occupy.  Default: all of them.  A worker that trains several members at
// once on one GPU lowers it (dtf_set_sm_limit) so that the members' kernels run side by side instead of
// queueing CTA by CTA; the value in force when a CUDA graph is captured is baked into that graph.
static int g_num_sms = 0, g_sm_limit = 0;
extern "C" int dtf_sm_count() {
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return (g_sm_limit > 0 && g_sm_limit < g_num_sms) ? g_sm_limit : g_num_sms;
}
extern "C" int dtf_stream_create(int device, cudaStream_t* out) {
    if (out == nullptr) return DTF_ERR_ARG;
    int prev = 0;
    if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return DTF_ERR_CUDA;
    const cudaError_t e = cudaStreamCreateWithFlags(out, cudaStreamNonBlocking);
    cudaSetDevice(prev);
    return e == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
extern "C" int dtf_stream_destroy(int device, cudaStream_t stream) {
    int prev = 0;
    if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) return DTF_ERR_CUDA;
    const cudaError_t e = cudaStreamDestroy(stream);
    cudaSetDevice(prev);
    return e == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
extern "C" int dtf_set_sm_limit(int n) {
    g_sm_limit = n > 0 ? n : 0;
    return dtf_sm_count();
}
static int num_sms() { return dtf_sm_count(); }

static int ilog2_exact(int v) {
    int s = 0;
    while ((1 << s) < v) ++s;
    return ((1 << s) == v) ? s : -1;
}

static int prep_gather(DtfGather* g) {
    if (g->dtype != DTF_BF16 && g->dtype != DTF_F32) return DTF_ERR_ARG;
    const int E = g->dtype == DTF_F32 ? 4 : 8;           // elements per 16-byte chunk
    if (g->C % E != 0) return DTF_ERR_SHAPE;
    const int cpt = g->C / E;
    g->cpt_shift = ilog2_exact(cpt);
    if (g->cpt_shift < 0) return DTF_ERR_SHAPE;
    if (g->KH * g->KW > 31) return DTF_ERR_SHAPE;
    if (g->C > dtf::kMaxC) return DTF_ERR_SHAPE;
    g->kchunks_real = g->KH * g->KW * cpt;
    return DTF_OK;
}

extern "C" int dtf_conv3x3_try(const DtfConvGemm* p, cudaStream_t stream);
extern "C" int dtf_conv3x3_wgrad_try(const DtfConvWgrad* p, cudaStream_t stream);
extern "C" int dtf_conv_s2_dgrad_try(const DtfConvGemm* p, cudaStream_t stream);

extern "C" int dtf_conv_gemm(DtfConvGemm* p, cudaStream_t stream) {
    using namespace dtf;
    int rc = prep_gather(&p->a);
    if (rc != DTF_OK) return rc;
    if (p->N % 16 != 0 || p->N < 16 || p->N > 256) return DTF_ERR_SHAPE;
    p->kchunks = (p->a.kchunks_real + 1) & ~1;
    const bool f32 = p->a.dtype == DTF_F32;
    if (!f32) {
        // 3x3 / stride 1: halo-tile kernel with shifted descriptor views (conv3x3_sm100.cu)
        rc = dtf_conv3x3_try(p, stream);
        if (rc != 0) return rc > 0 ? DTF_OK : rc;
        // stride-2 input gradient: parity-class kernel (conv_s2_sm100.cu)
        rc = dtf_conv_s2_dgrad_try(p, stream);
        if (rc != 0) return rc > 0 ? DTF_OK : rc;
    }
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
    if (!max_set) {
        if (cudaFuncSetAttribute(conv_gemm_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv_gemm_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        max_set = 1;
    }
    if (f32) dtf_launch(conv_gemm_kernel<4>, dim3(grid), dim3(kThreads), smem_bytes, stream, *p);
    else dtf_launch(conv_gemm_kernel<8>, dim3(grid), dim3(kThreads), smem_bytes, stream, *p);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

extern "C" int dtf_conv_wgrad(DtfConvWgrad* p, cudaStream_t stream) {
    using namespace dtf;
    int rc = prep_gather(&p->x);
    if (rc != DTF_OK) return rc;
    // gy side: 1x1 gather at the output pixel
    p->gy.KH = p->gy.KW = 1; p->gy.stride = 1; p->gy.pad = 0; p->gy.transposed = 0;
    p->gy.Hs = p->Ho; p->gy.Ws = p->Wo; p->gy.C = p->Cout; p->gy.dtype = p->x.dtype;
    rc = prep_gather(&p->gy);
    if (rc != DTF_OK) return rc;
    const bool f32 = p->x.dtype == DTF_F32;
    const int E = f32 ? 4 : 8;
    if (p->Cout % 8 != 0 || p->Cout > 256) return DTF_ERR_SHAPE;
    p->kchunks_real = p->x.kchunks_real;
    // UMMA N = chunks * E must be a multiple of 16: even chunk counts for bf16, multiples of four for fp32
    const int cmask = f32 ? 3 : 1;
    p->kchunks = (p->kchunks_real + cmask) & ~cmask;
    if (!f32) {
        rc = dtf_conv3x3_wgrad_try(p, stream);
        if (rc != 0) return rc > 0 ? DTF_OK : rc;
    }
    // columns per CTA: <= 24 chunks of bf16 (192 columns) / 16 chunks of fp32 (64 columns)
    const int per_max = f32 ? 16 : 24;
    int per = p->kchunks;
    int ny = 1;
    while (per > per_max) { ++ny; per = (((p->kchunks + ny - 1) / ny) + cmask) & ~cmask; }
    p->chunks_per_cta = per;
    ny = (p->kchunks + per - 1) / per;
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    const int a_off = (header + 1023) & ~1023;
    const int nz = (p->Cout + 127) / 128;              // 128 output channels per CTA (blockIdx.z)
    const int co_chunks = (p->Cout < 128 ? p->Cout : 128) / E;
    int buf_bytes, smem_bytes;
    if (!f32) {
        buf_bytes = (co_chunks + per) * kChunkBytes;
        // the A descriptor always spans 16 channel chunks: keep 32 KB of slack after the last buffer
        smem_bytes = a_off + 2 * buf_bytes + 16 * kChunkBytes;
    } else {
        // swizzled MN-major blocks of 32 channels / columns (16 KB each); the A descriptor spans four blocks
        const int a_blocks = (co_chunks + 7) / 8, b_blocks = (per + 7) / 8;
        buf_bytes = (a_blocks + b_blocks) * static_cast<int>(kSwBlockBytes);
        smem_bytes = a_off + 2 * buf_bytes + (4 - a_blocks) * static_cast<int>(kSwBlockBytes);
    }
    if (smem_bytes > 227 * 1024) return DTF_ERR_SMEM;
    p->smem_a_off = a_off;
    const int M = p->NB * p->Ho * p->Wo;
    const int tiles = (M + kTileM - 1) / kTileM;
    int gx = num_sms() / (ny * nz);
    if (gx < 1) gx = 1;
    if (gx > tiles) gx = tiles;
    static int set = 0;
    if (!set) {
        if (cudaFuncSetAttribute(conv_wgrad_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv_wgrad_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        set = 1;
    }
    dim3 grid(gx, ny, nz);
    if (f32) dtf_launch(conv_wgrad_kernel<4>, grid, dim3(kThreads), smem_bytes, stream, *p);
    else dtf_launch(conv_wgrad_kernel<8>, grid, dim3(kThreads), smem_bytes, stream, *p);
    return cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}

// Backward of one convolution in a single launch when both halves are 3x3 / stride 1 (see
// conv3x3_bwd_fused_kernel); otherwise two launches (dgrad, then wgrad).
extern "C" int dtf_conv3x3_bwd_fused_try(const DtfConvGemm* pd, const DtfConvWgrad* pw, cudaStream_t stream);

extern "C" int dtf_conv_bwd(DtfConvGemm* dgrad, DtfConvWgrad* wgrad, cudaStream_t stream) {
    using namespace dtf;
    int rc = prep_gather(&dgrad->a);
    if (rc != DTF_OK) return rc;
    dgrad->kchunks = (dgrad->a.kchunks_real + 1) & ~1;
    rc = prep_gather(&wgrad->x);
    if (rc != DTF_OK) return rc;
    wgrad->gy.KH = wgrad->gy.KW = 1; wgrad->gy.stride = 1; wgrad->gy.pad = 0; wgrad->gy.transposed = 0;
    wgrad->gy.Hs = wgrad->Ho; wgrad->gy.Ws = wgrad->Wo; wgrad->gy.C = wgrad->Cout; wgrad->gy.dtype = wgrad->x.dtype;
    rc = prep_gather(&wgrad->gy);
    if (rc != DTF_OK) return rc;
    if (dgrad->a.dtype != wgrad->x.dtype) return DTF_ERR_ARG;
    wgrad->kchunks_real = wgrad->x.kchunks_real;
    wgrad->kchunks = (wgrad->kchunks_real + 1) & ~1;
    if (wgrad->x.dtype == DTF_BF16) {
        rc = dtf_conv3x3_bwd_fused_try(dgrad, wgrad, stream);
        if (rc != 0) return rc > 0 ? DTF_OK : rc;
    }
    rc = dtf_conv_wgrad(wgrad, stream);
    if (rc != DTF_OK) return rc;
    return dtf_conv_gemm(dgrad, stream);
}
