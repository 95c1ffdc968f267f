// 3x3 / stride-1 / pad-1 convolutions (28 of ResNet-32's 34 convs, forward + dgrad + wgrad) on
// tcgen05 with a *halo tile* and *shifted descriptor views*.
//
// Idea: number the output pixels on a virtual grid  NB x (H+1) x (W+1)  in which one extra zero
// row per image and one extra zero column per row act as shared padding.  On that grid a 3x3 tap
// (dy, dx) is the constant flat shift  dy*(W+1) + dx.  A CTA therefore loads, for 128 consecutive
// virtual pixels, ONE halo tile of 128 + 2(W+1) + 2 input rows into shared memory — every input
// element read, BatchNorm+ReLU'd and bf16-packed once instead of nine times — stored in the
// no-swizzle UMMA layout [c/8][row][8].  The nine taps are then nine tcgen05.mma issues whose
// A-descriptor start address is the tile base advanced by (shift * 16 bytes); no im2col copy
// exists anywhere.  Virtual pixels that fall on the padding row/column are computed and dropped
// in the epilogue (6 % waste at W = 32, 21 % at W = 8).
//
// wgrad uses the same halo tile viewed MN-major (K = pixels): per 16-pixel K step one MMA per tap
// with N = Cin into TMEM columns [tap*Cin, (tap+1)*Cin), accumulating over all tiles of the CTA.
//
// Warp roles: 0-3 epilogue, 4 MMA issuer (+ TMEM alloc, weight bulk copy), 5-11 gather (one halo
// row per thread).  4 operand stages, 2 accumulators; two tiles' loads are in flight per gather
// thread before anything is stored, to cover L2 latency.
#include "conv_common.cuh"

#include <cstdlib>

namespace dtf {

// optional in-kernel timeline (CTA 0 only): compile with -DDTF_TIMING, read with dtf_debug_timeline
#ifdef DTF_TIMING
__device__ unsigned long long g_timeline[128];
#define DTF_MARK(slot) do { if (blockIdx.x == 0) g_timeline[slot] = globaltimer_ns(); } while (0)
#else
#define DTF_MARK(slot) do { } while (0)
#endif

// CTA layout of the halo kernels: 8 epilogue warps (two sets of four: a warp reads the TMEM lane quadrant
// warp % 4, so both sets see all 128 rows and split the COLUMNS of a tile), 1 MMA warp, 14 gather warps.
// In-kernel timelines of the previous layout (4 epilogue + 7 gather warps, one thread per halo row with all its
// channel chunks) showed the gather as the limiter at every width: 0.75 us per tile at 16 channels against 0.3 us
// of MMA issue and 0.25 us of epilogue, and one 2.1 - 2.8 us dependent chain per tile at 64 channels.  The work
// is now cut into (halo row, two-chunk) items spread over 448 gather threads, so the per-thread chain is the same
// two (or, with the second BN-backward source, four) 16-byte loads at every width.
#ifndef DTF_HALO_EPI_SETS
#define DTF_HALO_EPI_SETS 2
#endif
#ifndef DTF_HALO_GATHER_WARPS
#define DTF_HALO_GATHER_WARPS 14
#endif
constexpr int kHSets = DTF_HALO_EPI_SETS;                          // epilogue sets of four warps
constexpr int kHEpiWarps = 4 * kHSets;
constexpr int kHEpiThreads = kHEpiWarps * 32;                      // 256
constexpr int kHMmaWarp = kHEpiWarps;
constexpr int kHXWarp = kHMmaWarp + 1;                             // weight-gradient kernels: issues the im2col bulk copies
constexpr int kHGatherWarp0 = kHXWarp + 1;
constexpr int kHGatherWarps = DTF_HALO_GATHER_WARPS;
constexpr int kHThreads = (kHEpiWarps + 2 + kHGatherWarps) * 32;   // 768
constexpr int kHGatherThreads = kHGatherWarps * 32;               // 448
constexpr int kHMaxRows = 224;                                     // halo rows per tile the item maths assumes
constexpr int kHStages = 4;
constexpr int kHAcc = 4;     // TMEM accumulators of the forward / dgrad kernels: the MMA warp runs up to four tiles
                             // ahead of the epilogue sets (with two, accumulator hand-over was the critical cycle)

struct HaloGeom {
    FastDiv div_wp, div_hp;
    int H, W, Wp, Hp;
    int total;         // NB * Hp * Wp virtual pixels
    int num_tiles;
    int rows;          // halo rows per tile = 128 + 2*Wp + 2
    int cs;            // chunk stride in bytes (rows * 16 rounded up, plus 16: see fill_geom)
    int c8;            // input channel chunks
    int sign;          // +1 forward, -1 dgrad (gradient gather reads o - tap)
    int stage_bytes;
    int smem_b_off, smem_a_off;
    int tap0, ntaps;   // wgrad: taps handled by blockIdx.y == 0; the rest by y == 1
    int gy_c8;         // wgrad: output-channel chunks
    int bulk;          // wgrad: im2col operand built by per-tap shared -> shared bulk copies (chunk stride = cs)
};

struct __align__(16) HaloHeader {
    uint64_t full[kHStages];
    uint64_t empty[kHStages];
    uint64_t halo[kHStages];     // weight gradient: the x halo tile of a stage is in shared memory
    uint64_t acc_full[4];
    uint64_t acc_empty[4];
    uint64_t bfull;
    uint32_t tmem_base;
    uint32_t pad;
};

// virtual flat index -> dense pixel index; false for padding positions
DTF_DEVINL bool decode_virtual(int q, const HaloGeom& G, uint32_t& pix) {
    if (q < 0 || q >= G.total) return false;
    const uint32_t R = fdiv(static_cast<uint32_t>(q), G.div_wp);
    const uint32_t xx = static_cast<uint32_t>(q) - R * G.Wp;
    const uint32_t n = fdiv(R, G.div_hp);
    const uint32_t yy = R - n * G.Hp;
    if (xx >= static_cast<uint32_t>(G.W) || yy >= static_cast<uint32_t>(G.H)) return false;
    pix = (n * G.H + yy) * G.W + xx;
    return true;
}

// prologue on one 16-byte chunk (8 channels starting at cb)
DTF_DEVINL uint4 apply_prologue(int mode, const uint4& raw, const uint4& raw2, const float* p0, const float* p1,
                                const float* p2, int cb) {
    if (mode == DTF_PRO_NONE) return raw;
    float f[8];
    unpack8(raw, f);
    const float4 a0 = *reinterpret_cast<const float4*>(p0 + cb), a1 = *reinterpret_cast<const float4*>(p0 + cb + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(p1 + cb), b1 = *reinterpret_cast<const float4*>(p1 + cb + 4);
    const float s[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    const float t[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    if (mode == DTF_PRO_BN_RELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaxf(fmaf(f[i], s[i], t[i]), 0.f);
    } else {
        float x[8];
        unpack8(raw2, x);
        const float4 c0 = *reinterpret_cast<const float4*>(p2 + cb), c1 = *reinterpret_cast<const float4*>(p2 + cb + 4);
        const float u[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = fmaf(f[i], s[i], fmaf(x[i], t[i], u[i]));
    }
    return pack8(f);
}

// One gather item: two consecutive 16-byte chunks (16 channels) of one row.
struct Item {
    const dtf_bf16* src;   // first channel of the item in the source tensor
    const dtf_bf16* src2;  // second source (BN backward) or nullptr
    uint32_t dst;          // shared address of the item's first chunk
    int cb;                // first channel
    bool live;             // this thread has an item in this slot
    bool valid;            // the row maps to a real pixel (otherwise zeros are stored)
};

// Load + prologue + store NI items: every 16-byte load is issued before the first one is consumed.
template <int NI>
DTF_DEVINL void run_items(const Item (&it)[NI], int cs, int mode, const float* p0, const float* p1, const float* p2) {
    uint4 a[NI][2], b[NI][2];
#pragma unroll
    for (int u = 0; u < NI; ++u)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            a[u][c] = make_uint4(0, 0, 0, 0);
            b[u][c] = a[u][c];
            if (it[u].live && it[u].valid) {
                a[u][c] = ldg16(it[u].src + c * 8);
                if (mode == DTF_PRO_BN_BWD) b[u][c] = ldg16(it[u].src2 + c * 8);
            }
        }
#pragma unroll
    for (int u = 0; u < NI; ++u)
#pragma unroll
        for (int c = 0; c < 2; ++c)
            if (it[u].live) {
                const uint4 v = it[u].valid ? apply_prologue(mode, a[u][c], b[u][c], p0, p1, p2, it[u].cb + c * 8)
                                            : make_uint4(0, 0, 0, 0);
                sts16(it[u].dst + c * cs, v);
            }
}

// =========================================================================================
// forward / dgrad
// =========================================================================================
// The body is a device function parameterised by (cta, ncta) so that the same code runs as a
// stand-alone kernel and as one half of the fused backward launch (dgrad CTAs + wgrad CTAs).
template <int C8>
DTF_DEVINL void conv3x3_body(const DtfConvGemm& P, const HaloGeom& G, const int cta, const int ncta, uint8_t* smem) {
    constexpr int kPairs = C8 / 2;            // UMMA K steps (16 channels) per tap = gather items per halo row
    constexpr int kJob = (C8 == 2 && 2 * kHMaxRows <= kHGatherThreads) ? 2 : 1;   // tiles gathered together
    constexpr int kNI = (kJob * kHMaxRows * kPairs + kHGatherThreads - 1) / kHGatherThreads;   // items per gather thread per job
    HaloHeader* hdr = reinterpret_cast<HaloHeader*>(smem);
    SmemConsts* cst = reinterpret_cast<SmemConsts*>(smem + 1024);
    uint8_t* smem_b = smem + G.smem_b_off;
    uint8_t* smem_a = smem + G.smem_a_off;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = P.N;
    const uint32_t acc_cols = tmem_cols_pow2(2 * N);
    const int my_tiles = (G.num_tiles - cta + ncta - 1) / ncta;
    // epilogue sets: with >= 32 output channels both sets work on every tile (half of the columns each)
    constexpr int kGroups = C8 / 2;                               // 16-column groups of a tile
    // Two epilogue sets: with >= 32 output channels both work on every tile (half of the 16-column groups each);
    // with 16 channels (one group) they take alternate tiles, each with its own accumulator, so that a set's
    // inter-tile work (decode, auxiliary prefetch, statistics) overlaps the other set's tile.
    constexpr bool kSplit = kHSets == 2 && kGroups >= 2;
    constexpr bool kAlt = kHSets == 2 && kGroups == 1;
    constexpr int kGPS = kSplit ? kGroups / 2 : kGroups;          // groups per epilogue set
    constexpr int kEpiPerAcc = kSplit ? kHEpiThreads : 128;       // threads that release one accumulator
    constexpr int kEpiActive = kHSets == 2 ? kHEpiThreads : 128;  // threads in the epilogue role

    if (tid == 0) {
        for (int i = 0; i < kHStages; ++i) {
            mbar_init(smem_u32(&hdr->full[i]), kHGatherThreads);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
        }
        for (int i = 0; i < kHAcc; ++i) {
            mbar_init(smem_u32(&hdr->acc_full[i]), 1);
            mbar_init(smem_u32(&hdr->acc_empty[i]), kEpiPerAcc);
        }
        mbar_init(smem_u32(&hdr->bfull), 1);
        mbar_fence_init();
    }
    for (int c = tid; c < 2 * kMaxC + 8 * 2 * 64; c += kHThreads) (&cst->bins[0][0])[c] = 0.f;   // bins + wbins
    if (warp == kHMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    griddep_launch_dependents();
    griddep_wait();
    fill_prologue(P.a, cst->p0, cst->p1, cst->p2, tid, kHThreads);
    if (P.epi_mode == DTF_EPI_RELU_MASK)
        for (int c = tid; c < N; c += kHThreads)
            bn_fwd_consts(P.epi_bn, c, cst->e_scale[c], cst->e_shift[c], cst->e_mean[c], cst->e_invstd[c]);
    if (cta == 0 && P.a.update_running && P.a.mode == DTF_PRO_BN_RELU && !P.a.bn.use_running)
        update_running(P.a.bn, tid, kHThreads);
    if (tid == 0) DTF_MARK(0);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    if (tid == 0) DTF_MARK(1);

    if (warp == kHMmaWarp) {
        // Warp-uniform loop: every lane computes the (uniform) descriptors, one elected lane issues.
        // (Issuing from a divergent `lane == 0` branch makes ptxas serialise every operand through
        // R2UR waterfall loops: ~170 cycles per MMA instead of ~20.)
        const uint32_t wbytes = static_cast<uint32_t>(P.kchunks) * N * 16u;
        if (elect_one()) {
            mbar_arrive_expect_tx(smem_u32(&hdr->bfull), wbytes);
            bulk_g2s(smem_u32(smem_b), P.wpk, wbytes, smem_u32(&hdr->bfull));
        }
        __syncwarp();
        mbar_wait(smem_u32(&hdr->bfull), 0, 0x31);
        DTF_MARK(2);
        const uint32_t idesc = umma_idesc_bf16(kTileM, N, false, false);
        // Descriptor arithmetic hoisted out of the tile loop: per MMA two 32-bit adds and two multiply-adds.
        // (A descriptor low word = start address >> 4; shifting the view by `s` halo rows adds `s`.)
        uint32_t a_tap[9];
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int dy = tap / 3 - 1, dx = tap % 3 - 1;
            a_tap[tap] = static_cast<uint32_t>((1 + G.sign * dy) * G.Wp + (1 + G.sign * dx));   // halo row of output row 0
        }
        const uint32_t a_pair = static_cast<uint32_t>((2 * G.cs) >> 4);          // two chunks further in the halo tile
        const uint32_t b_pair = static_cast<uint32_t>(2 * N), b_tap = static_cast<uint32_t>(C8 * N);
        const uint64_t adesc0 = umma_desc(smem_u32(smem_a), G.cs, 128);
        const uint64_t bdesc0 = umma_desc(smem_u32(smem_b), N * 16, 128);
        // the start-address field (bits 0-13 of the low word, units of 16 bytes) never carries out of the field for
        // addresses inside the 227 KB of shared memory, so the shifted views only touch the low words
        const uint32_t a_lo0 = static_cast<uint32_t>(adesc0), a_hi = static_cast<uint32_t>(adesc0 >> 32);
        const uint32_t b_lo0 = static_cast<uint32_t>(bdesc0), b_hi = static_cast<uint32_t>(bdesc0 >> 32);
        const uint32_t stage_units = static_cast<uint32_t>(G.stage_bytes >> 4);
        for (int it = 0; it < my_tiles; ++it) {
            const uint32_t a = it % kHAcc, st = it % kHStages;
            mbar_wait(smem_u32(&hdr->acc_empty[a]), ((it / kHAcc) & 1) ^ 1, 0x32);
            if (it < 8) DTF_MARK(88 + it);
            mbar_wait(smem_u32(&hdr->full[st]), (it / kHStages) & 1, 0x33);
            tc_fence_after_sync();
            if (it < 4) DTF_MARK(8 + it);
            if (it < 8) DTF_MARK(96 + it);
            const uint32_t a_lo_st = a_lo0 + st * stage_units;
            const uint32_t d_tmem = tmem_base + a * N;
            if (elect_one()) {
                umma_bf16_lohi<false>(d_tmem, a_lo_st + a_tap[0], a_hi, b_lo0, b_hi, idesc);
#pragma unroll
                for (int tap = 0; tap < 9; ++tap)
#pragma unroll
                    for (int kp = 0; kp < kPairs; ++kp)
                        if ((tap | kp) != 0)
                            umma_bf16_lohi<true>(d_tmem, a_lo_st + (a_tap[tap] + kp * a_pair), a_hi,
                                                 b_lo0 + (tap * b_tap + kp * b_pair), b_hi, idesc);
                umma_commit(smem_u32(&hdr->empty[st]));
                umma_commit(smem_u32(&hdr->acc_full[a]));
            }
            __syncwarp();
            if (it < 4) DTF_MARK(12 + it);
            if (it < 8) DTF_MARK(104 + it);
        }
    } else if (warp >= kHGatherWarp0) {
        // ================= gather: item = (halo row, pair of chunks), spread over all gather threads =================
        const int g = tid - kHGatherWarp0 * 32;
        const int per_tile = G.rows * kPairs;
        for (int it = 0; it < my_tiles; it += kJob) {
            const int nt = min(kJob, my_tiles - it);
            Item items[kNI];
#pragma unroll
            for (int u = 0; u < kNI; ++u) {
                int idx = g + u * kHGatherThreads;
                int tl = 0;
                if (kJob == 2 && idx >= per_tile) { idx -= per_tile; tl = 1; }
                const bool live = tl < nt && idx < per_tile;
                const int row = kPairs == 1 ? idx : (kPairs == 2 ? idx >> 1 : idx >> 2);
                const int part = idx - row * kPairs;
                uint32_t pix = 0;
                const int tile = cta + (it + tl) * ncta;
                items[u].live = live;
                items[u].valid = live && decode_virtual(tile * kTileM - G.Wp - 1 + row, G, pix);
                const size_t off = static_cast<size_t>(pix) * (C8 * 8) + part * 16;
                items[u].src = P.a.src + off;
                items[u].src2 = P.a.src2 ? P.a.src2 + off : nullptr;
                items[u].cb = part * 16;
                items[u].dst = smem_u32(smem_a) + ((it + tl) % kHStages) * G.stage_bytes + (2 * part) * G.cs + row * 16;
            }
            if (g == 0 && it < 8) DTF_MARK(16 + it);
#pragma unroll
            for (int u = 0; u < kJob; ++u)
                if (u < nt) mbar_wait(smem_u32(&hdr->empty[(it + u) % kHStages]), (((it + u) / kHStages) & 1) ^ 1, 0x34);
            if (g == 0 && it < 8) DTF_MARK(64 + it);
            run_items<kNI>(items, G.cs, P.a.mode, cst->p0, cst->p1, cst->p2);
            if (g == 0 && it < 8) DTF_MARK(72 + it);
            fence_proxy_async_smem();
            if (g == 0 && it < 8) DTF_MARK(80 + it);
#pragma unroll
            for (int u = 0; u < kJob; ++u)
                if (u < nt) mbar_arrive(smem_u32(&hdr->full[(it + u) % kHStages]));
            if (g == 0 && it < 8) DTF_MARK(24 + it);
        }
    } else if (warp < kHMmaWarp && (kHSets == 2 || warp < 4)) {
        // ================= epilogue (N == C8 * 8, compile-time): set = warp / 4 owns kGPS column groups =================
        const int set = warp >> 2, quad = warp & 3;
        const int g0 = kSplit ? set * kGPS : 0;            // first 16-column group of this set
        // BN statistics in registers across tiles: only in the one-set layout (96 registers per thread); with two
        // sets (80 registers) they go through the warp butterfly + shared bins every tile
        constexpr bool kRegStats = kHSets == 1 && kGPS * 16 <= 32;
        const int t_first = kAlt ? set : 0;
        constexpr int t_step = kAlt ? 2 : 1;
        const dtf_bf16* aux_src = epilogue_aux_ptr(P);
        float* wb = &cst->wbins[warp][0][0];                 // this warp's private statistics bins
        float acc1[kRegStats ? kGPS * 16 : 1], acc2[kRegStats ? kGPS * 16 : 1];
#pragma unroll
        for (int i = 0; i < (kRegStats ? kGPS * 16 : 1); ++i) { acc1[i] = 0.f; acc2[i] = 0.f; }
        // the auxiliary tensor (residual / mask source) does not depend on the MMA: fetch it one tile ahead
        uint4 aux_nxt[2 * kGPS];
        uint32_t pix_nxt = 0;
        bool valid_nxt = false;
        if (my_tiles > t_first) valid_nxt = decode_virtual((cta + t_first * ncta) * kTileM + quad * 32 + lane, G, pix_nxt);
#pragma unroll
        for (int j = 0; j < 2 * kGPS; ++j)
            aux_nxt[j] = (aux_src != nullptr && valid_nxt)
                             ? ldg16(aux_src + static_cast<size_t>(pix_nxt) * N + g0 * 16 + j * 8)
                             : make_uint4(0, 0, 0, 0);
        for (int it = t_first; it < my_tiles; it += t_step) {
            const uint32_t a = it % kHAcc;
            const bool valid = valid_nxt;
            const size_t obase = static_cast<size_t>(pix_nxt) * N;
            uint4 aux[2 * kGPS];
#pragma unroll
            for (int j = 0; j < 2 * kGPS; ++j) aux[j] = aux_nxt[j];
            if (it + t_step < my_tiles) {
                const int tile_n = cta + (it + t_step) * ncta;
                valid_nxt = decode_virtual(tile_n * kTileM + quad * 32 + lane, G, pix_nxt);
#pragma unroll
                for (int j = 0; j < 2 * kGPS; ++j)
                    aux_nxt[j] = (aux_src != nullptr && valid_nxt)
                                     ? ldg16(aux_src + static_cast<size_t>(pix_nxt) * N + g0 * 16 + j * 8)
                                     : make_uint4(0, 0, 0, 0);
            }
            if (tid == 0 && it < 8) DTF_MARK(112 + it);
            mbar_wait(smem_u32(&hdr->acc_full[a]), (it / kHAcc) & 1, 0x35);
            tc_fence_after_sync();
            if (tid == 0 && it < 8) DTF_MARK(32 + it);
#pragma unroll
            for (int gi = 0; gi < kGPS; ++gi) {
                const int cg = (g0 + gi) * 16;
                float v[16];
                tmem_ld16(tmem_base + ((quad * 32u) << 16) + a * N + cg, v);
                epilogue_group(P, cst, v, valid, obase, cg, lane, &aux[2 * gi],
                               kRegStats ? &acc1[gi * 16] : nullptr, kRegStats ? &acc2[gi * 16] : nullptr, wb);
            }
            tc_fence_before_sync();
            mbar_arrive(smem_u32(&hdr->acc_empty[a]));
            if (tid == 0 && it < 8) DTF_MARK(40 + it);
        }
        if (kRegStats && P.epi_stats_out) {
#pragma unroll
            for (int gi = 0; gi < kGPS; ++gi) flush_stats16(cst, &acc1[gi * 16], &acc2[gi * 16], (g0 + gi) * 16, lane, wb);
        }
        if (P.epi_stats_out) {
            named_bar_sync(1, kEpiActive);
            for (int c = tid; c < 2 * N; c += kEpiActive) {      // c < N: sums, c >= N: second statistic
                const int k = c >= N ? 1 : 0, ch = c - k * N;
                float tot = 0.f;
#pragma unroll
                for (int w = 0; w < kHEpiWarps; ++w) tot += cst->wbins[w][k][ch];
                atomicAdd(P.epi_stats_out + c, tot);
            }
        }
    }
    if (tid == 0) DTF_MARK(3);
    tc_fence_before_sync();
    __syncthreads();
    if (tid == 0) DTF_MARK(4);
    if (warp == kHMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, acc_cols);
    }
}

template <int C8>
__global__ void __launch_bounds__(kHThreads, 1)
conv3x3_kernel(const __grid_constant__ DtfConvGemm P, const __grid_constant__ HaloGeom G) {
    extern __shared__ __align__(1024) uint8_t smem[];
    conv3x3_body<C8>(P, G, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x), smem);
}

// =========================================================================================
// wgrad:  dw[co][tap*Cin + ci] += sum_pixels gy'[p][co] * act(x)[p + shift(tap)][ci]
//
// Per 128-pixel tile the gather warps (1) load the x halo tile into shared memory with its prologue applied
// once, (2) expand the halo into an explicit im2col operand [tap*C8 + c8][128 rows][8] with shared->shared
// 16-byte copies (the nine shifted views of the halo), so that (3) one tcgen05.mma per 16-pixel K step covers
// ALL taps at once (N = 9*Cin, split into groups of <= 256 columns) instead of one MMA per tap; the epilogue
// warps load the gy' tile meanwhile.  Both operands are MN-major views (K = pixels).  The accumulator
// D[co][tap*Cin+ci] lives in TMEM across all tiles of the CTA and is flushed once with vector reductions
// (red.global.add.v4.f32), the two epilogue sets taking alternate 16-column groups.
// =========================================================================================
DTF_DEVINL uint4 lds16(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
DTF_DEVINL void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// x side of one weight-gradient tile: halo load (prologue applied) -> arrive on the stage's halo barrier.
// GT threads cooperate on a tile (all 448 gather threads, or one group of 224 when two groups alternate tiles);
// `gl` is the thread's index inside its group, the group handles tiles t_first, t_first + t_step, ...
template <int C8, int GT>
DTF_DEVINL void wgrad_gather(const DtfConvWgrad& P, const HaloGeom& G, HaloHeader* hdr, SmemConsts* cst,
                             uint8_t* smem_a, int cta, int ncta, int my_tiles, int nstages,
                             const uint32_t* ex_tab, int nchunks, int gy_bytes, int im_bytes, int gl, int t_first,
                             int t_step, int bar_id) {
    constexpr int kPairs = C8 / 2;
    constexpr int Cin = C8 * 8;
    constexpr int kNI = (kHMaxRows * kPairs + GT - 1) / GT;
    const int per_tile = G.rows * kPairs;
    for (int it = t_first; it < my_tiles; it += t_step) {
        const int tile = cta + it * ncta;
        const uint32_t st = it % nstages;
        const uint32_t gy_base = smem_u32(smem_a) + st * G.stage_bytes;
        const uint32_t im_base = gy_base + gy_bytes;
        const uint32_t halo_base = im_base + im_bytes;
        Item items[kNI];
#pragma unroll
        for (int u = 0; u < kNI; ++u) {
            const int idx = gl + u * GT;
            const bool live = idx < per_tile;
            const int row = kPairs == 1 ? idx : (kPairs == 2 ? idx >> 1 : idx >> 2);
            const int part = idx - row * kPairs;
            uint32_t pix = 0;
            items[u].live = live;
            items[u].valid = live && decode_virtual(tile * kTileM - G.Wp - 1 + row, G, pix);
            items[u].src = P.x.src + static_cast<size_t>(pix) * Cin + part * 16;
            items[u].src2 = nullptr;
            items[u].cb = part * 16;
            items[u].dst = halo_base + (2 * part) * G.cs + row * 16;
        }
        mbar_wait(smem_u32(&hdr->empty[st]), ((it / nstages) & 1) ^ 1, 0x42);
        if (gl == 0 && it < 4) DTF_MARK(16 + 4 * it);
        run_items<kNI>(items, G.cs, P.x.mode, cst->p0, cst->p1, cst->p2);
        if (gl == 0 && it < 4) DTF_MARK(17 + 4 * it);
        if (G.bulk) {
            // hand the halo tile to the expander warp: it builds the im2col operand with one shared -> shared bulk
            // copy per tap (the whole halo tile shifted by the tap's row offset)
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->halo[st]));
            continue;
        }
        named_bar_sync(bar_id, GT);                   // the halo tile is complete (this group's warps only)
        if (gl == 0 && it < 4) DTF_MARK(18 + 4 * it);
        // im2col expansion by the gather threads: chunk (tap, c8), row r  <-  halo chunk c8, row r + shift(tap);
        // source offsets per im2col chunk come from a table built once per CTA (ex_tab); two copies in flight
        const int nitems = nchunks * kTileM;
        int i = gl;
        for (; i + GT < nitems; i += 2 * GT) {
            const int i2 = i + GT;
            const uint32_t r0 = (i & (kTileM - 1)) * 16, r1 = (i2 & (kTileM - 1)) * 16;
            const int j0 = i >> 7, j1 = i2 >> 7;
            const uint4 v0 = lds16(halo_base + ex_tab[j0] + r0);
            const uint4 v1 = lds16(halo_base + ex_tab[j1] + r1);
            sts16(im_base + j0 * kChunkBytes + r0, v0);
            sts16(im_base + j1 * kChunkBytes + r1, v1);
        }
        if (i < nitems) {
            const uint32_t r0 = (i & (kTileM - 1)) * 16;
            const int j0 = i >> 7;
            sts16(im_base + j0 * kChunkBytes + r0, lds16(halo_base + ex_tab[j0] + r0));
        }
        fence_proxy_async_smem();
        mbar_arrive(smem_u32(&hdr->full[st]));
        if (gl == 0 && it < 4) DTF_MARK(19 + 4 * it);
    }
}

template <int C8>
DTF_DEVINL void conv3x3_wgrad_body(const DtfConvWgrad& P, const HaloGeom& G, const int cta, const int ncta,
                                   const int cta_y, uint8_t* smem) {
    HaloHeader* hdr = reinterpret_cast<HaloHeader*>(smem);
    SmemConsts* cst = reinterpret_cast<SmemConsts*>(smem + 1024);
    uint8_t* smem_a = smem + G.smem_a_off;   // per stage: [gy: gy_c8 chunks][im2col: ntaps*C8 chunks][halo: C8 x rows]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int Cin = C8 * 8;
    const int tap_begin = cta_y == 0 ? 0 : G.ntaps;
    const int ntaps = cta_y == 0 ? G.ntaps : 9 - G.ntaps;
    const int nchunks = ntaps * C8;                          // im2col chunks (8 columns each)
    const int ngroups = (nchunks + 31) / 32;                 // MMA column groups of <= 256
    const int gchunks = (((nchunks + ngroups - 1) / ngroups) + 1) & ~1;
    const uint32_t acc_cols = tmem_cols_pow2(nchunks * 8);
    const int my_tiles = (G.num_tiles - cta + ncta - 1) / ncta;
    const int gy_bytes = G.gy_c8 * kChunkBytes;
    const int im_bytes = G.tap0;                             // byte size of the im2col area (set by the host)
    const int nstages = G.sign;                              // number of operand stages (set by the host)
    const bool walt = nstages >= 2;                          // gather groups alternate tiles

    if (tid == 0) {
        for (int i = 0; i < kHStages; ++i) {
            // operands of a stage: the im2col operand (bulk copies: one expect_tx arrival + their bytes; otherwise the
            // gather threads that expand it) + gy' loaders
            mbar_init(smem_u32(&hdr->full[i]), (G.bulk ? 1 : (walt ? kHGatherThreads / 2 : kHGatherThreads)) + kHEpiThreads);
            mbar_init(smem_u32(&hdr->empty[i]), 1);
            // x gather warps (one group of seven when the groups alternate)
            mbar_init(smem_u32(&hdr->halo[i]), walt ? kHGatherThreads / 2 : kHGatherThreads);
        }
        mbar_init(smem_u32(&hdr->acc_full[0]), 1);
        mbar_fence_init();
    }
    if (warp == kHMmaWarp) tmem_alloc(smem_u32(&hdr->tmem_base), acc_cols);
    griddep_launch_dependents();
    griddep_wait();
    fill_prologue(P.x, cst->p0, cst->p1, cst->p2, tid, kHThreads);
    fill_prologue(P.gy, cst->q0, cst->q1, cst->q2, tid, kHThreads);
    // im2col expansion table: chunk j = (tap t, channel chunk c8) reads halo chunk c8 shifted by the tap
    uint32_t* ex_tab = reinterpret_cast<uint32_t*>(&cst->bins[0][0]);     // the statistics bins are unused here
    for (int j = tid; j < nchunks; j += kHThreads) {
        const int t = j / C8, c8 = j - t * C8;
        const int tap = tap_begin + t;
        const int shift = (tap / 3) * G.Wp + (tap % 3);                   // (1+dy)*Wp + (1+dx)
        ex_tab[j] = static_cast<uint32_t>(c8 * G.cs + shift * 16);
    }
    if (tid == 0) DTF_MARK(0);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = hdr->tmem_base;
    if (tid == 0) DTF_MARK(1);

    if (warp == kHMmaWarp) {
        if (my_tiles > 0) {          // warp-uniform loop, one elected lane issues
            const uint64_t gdesc0 = umma_desc(smem_u32(smem_a), 128, kChunkBytes);
            const uint32_t bstride = G.bulk ? static_cast<uint32_t>(G.cs) : static_cast<uint32_t>(kChunkBytes);   // im2col chunk stride
            const uint64_t idesc0 = umma_desc(smem_u32(smem_a) + gy_bytes, 128, bstride);
            const uint32_t stage_units = static_cast<uint32_t>(G.stage_bytes >> 4);
            for (int it = 0; it < my_tiles; ++it) {
                const uint32_t st = it % nstages;
                mbar_wait(smem_u32(&hdr->full[st]), (it / nstages) & 1, 0x41);
                tc_fence_after_sync();
                if (elect_one()) {
                    if (it < 4) DTF_MARK(8 + it);
                    for (int grp = 0; grp < ngroups; ++grp) {
                        const int c0 = grp * gchunks;
                        const int nc = min(gchunks, nchunks - c0);
                        const uint32_t idesc = umma_idesc_bf16(kTileM, nc * 8, true, true);
#pragma unroll
                        for (int ks = 0; ks < kTileM / 16; ++ks)
                            umma_bf16(tmem_base + c0 * 8, gdesc0 + st * stage_units + ks * 16,
                                      idesc0 + st * stage_units + c0 * (bstride >> 4) + ks * 16, idesc, (it | ks) != 0);
                    }
                    umma_commit(smem_u32(&hdr->empty[st]));
                    if (it < 4) DTF_MARK(12 + it);
                }
                __syncwarp();
            }
            if (elect_one()) umma_commit(smem_u32(&hdr->acc_full[0]));
        }
        __syncwarp();
    } else if (warp == kHXWarp) {
        // ================= expander: im2col operand by the async copy engine =================
        // Tap t of the operand is the whole halo tile [C8][rows] moved up by shift(t) rows, so one bulk copy of
        // (C8 - 1) * cs + 2 KB bytes per tap builds it with the halo's own chunk stride (issuing a bulk copy costs
        // ~40 ns whatever its size: 2 KB copies per (tap, chunk) were slower than the threads they replaced).
        if (G.bulk) {
            const uint32_t tap_bytes = static_cast<uint32_t>((C8 - 1) * G.cs + kChunkBytes);
            for (int it = 0; it < my_tiles; ++it) {
                const uint32_t st = it % nstages;
                const uint32_t im_base = smem_u32(smem_a) + st * G.stage_bytes + gy_bytes;
                const uint32_t halo_base = im_base + im_bytes;
                mbar_wait(smem_u32(&hdr->halo[st]), (it / nstages) & 1, 0x45);
                if (it < 4) DTF_MARK(64 + it);
                if (elect_one()) {
                    mbar_arrive_expect_tx(smem_u32(&hdr->full[st]), static_cast<uint32_t>(ntaps) * tap_bytes);
                    for (int t = 0; t < ntaps; ++t) {
                        const int tap = tap_begin + t;
                        const uint32_t shift = static_cast<uint32_t>((tap / 3) * G.Wp + (tap % 3));
                        bulk_s2s(im_base + t * (C8 * G.cs), halo_base + shift * 16, tap_bytes, smem_u32(&hdr->full[st]));
                    }
                }
                __syncwarp();
                if (it < 4) DTF_MARK(72 + it);
            }
        }
    } else if (warp >= kHGatherWarp0) {
        const int g = tid - kHGatherWarp0 * 32;
        if (walt) {      // two groups of seven warps, alternate tiles, each with its own operand stage in flight
            const int grp = g >= kHGatherThreads / 2 ? 1 : 0;
            wgrad_gather<C8, kHGatherThreads / 2>(P, G, hdr, cst, smem_a, cta, ncta, my_tiles, nstages, ex_tab, nchunks,
                                                  gy_bytes, im_bytes, g - grp * (kHGatherThreads / 2), grp, 2, 2 + grp);
        } else {         // one operand stage only (64 channels): all gather warps work on the same tile
            wgrad_gather<C8, kHGatherThreads>(P, G, hdr, cst, smem_a, cta, ncta, my_tiles, nstages, ex_tab, nchunks,
                                              gy_bytes, im_bytes, g, 0, 1, 2);
        }
    } else {
        // The eight epilogue warps have nothing to flush until the last tile: they load the gy' tile (128 rows, items
        // of two chunks, BN-backward prologue with its two sources) while the gather warps load and expand the x
        // halo, so the two halves of a tile's operand traffic run side by side instead of back to back.
        const int gy_pairs = G.gy_c8 >> 1;
        const int gy_items = kTileM * gy_pairs;
        for (int it = 0; it < my_tiles; ++it) {
            const int tile = cta + it * ncta;
            const uint32_t st = it % nstages;
            mbar_wait(smem_u32(&hdr->empty[st]), ((it / nstages) & 1) ^ 1, 0x44);
            for (int idx = tid; idx < gy_items; idx += kHEpiThreads) {
                const int row = gy_pairs == 1 ? idx : (gy_pairs == 2 ? idx >> 1 : idx >> 2);
                const int part = idx - row * gy_pairs;
                uint32_t pix2 = 0;
                Item jg[1];
                jg[0].live = true;
                jg[0].valid = decode_virtual(tile * kTileM + row, G, pix2);
                const size_t off = static_cast<size_t>(pix2) * P.Cout + part * 16;
                jg[0].src = P.gy.src + off;
                jg[0].src2 = P.gy.src2 ? P.gy.src2 + off : nullptr;
                jg[0].cb = part * 16;
                jg[0].dst = smem_u32(smem_a) + st * G.stage_bytes + (2 * part) * kChunkBytes + row * 16;
                run_items<1>(jg, kChunkBytes, P.gy.mode, cst->q0, cst->q1, cst->q2);
            }
            fence_proxy_async_smem();
            mbar_arrive(smem_u32(&hdr->full[st]));
        }
        if (my_tiles > 0) {
            mbar_wait(smem_u32(&hdr->acc_full[0]), 0, 0x43);
            tc_fence_after_sync();
            if (tid == 0) DTF_MARK(32);
            const int set = warp >> 2, quad = warp & 3;
            const int co = quad * 32 + lane;
            const int ktot = 9 * Cin;
            const int kbase = tap_begin * Cin;
            for (int cg = set * 16; cg < nchunks * 8; cg += 16 * kHSets) {     // the sets take alternate 16-column groups
                float v[16];
                tmem_ld16(tmem_base + ((quad * 32u) << 16) + cg, v);
                if (co < P.Cout) {
                    float* dst = P.dw + static_cast<size_t>(co) * ktot + kbase + cg;
#pragma unroll
                    for (int i = 0; i < 16; i += 4) red_add_v4(dst + i, v[i], v[i + 1], v[i + 2], v[i + 3]);
                }
            }
            if (tid == 0) DTF_MARK(33);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (tid == 0) DTF_MARK(4);
    if (warp == kHMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, acc_cols);
    }
}

template <int C8>
__global__ void __launch_bounds__(kHThreads, 1)
conv3x3_wgrad_kernel(const __grid_constant__ DtfConvWgrad P, const __grid_constant__ HaloGeom G) {
    extern __shared__ __align__(1024) uint8_t smem[];
    conv3x3_wgrad_body<C8>(P, G, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x), static_cast<int>(blockIdx.y), smem);
}

// One launch for the two backward GEMMs of a 3x3 convolution: CTAs [0, n_dgrad) compute the input
// gradient, the remaining CTAs the weight gradient (linear index -> (x, y) of the wgrad grid).  The two
// are independent (both only read gy'), so sharing a launch removes one launch latency per layer and
// lets the two halves fill the SMs that either alone would leave idle in the small-tile stages.
template <int C8>
__global__ void __launch_bounds__(kHThreads, 1)
conv3x3_bwd_fused_kernel(const __grid_constant__ DtfConvGemm Pd, const __grid_constant__ HaloGeom Gd,
                         const __grid_constant__ DtfConvWgrad Pw, const __grid_constant__ HaloGeom Gw,
                         const int n_dgrad, const int wgrad_gx) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int b = static_cast<int>(blockIdx.x);
    if (b < n_dgrad) {
        conv3x3_body<C8>(Pd, Gd, b, n_dgrad, smem);
    } else {
        const int w = b - n_dgrad;
        conv3x3_wgrad_body<C8>(Pw, Gw, w % wgrad_gx, wgrad_gx, w / wgrad_gx, smem);
    }
}

}  // namespace dtf

// =========================================================================================
// host side: eligibility + geometry + launch.  Called by dtf_conv_gemm / dtf_conv_wgrad.
// =========================================================================================
static dtf::FastDiv make_fastdiv(uint32_t d) {
    dtf::FastDiv f;
    f.d = d;
    uint32_t s = 0;
    while ((1u << s) < d) ++s;
    f.s = s;
    f.m = static_cast<uint32_t>(((1ull << 32) * ((1ull << s) - d)) / d + 1);
    return f;
}

extern "C" int dtf_sm_count();
static int halo_sms() { return dtf_sm_count(); }

static bool halo_disabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("DTF_CONV_GENERIC");
        v = (e && e[0] == '1') ? 1 : 0;
    }
    return v == 1;
}

static bool fill_geom(dtf::HaloGeom* G, int NB, int H, int W, int C) {
    G->H = H; G->W = W; G->Wp = W + 1; G->Hp = H + 1;
    G->div_wp = make_fastdiv(G->Wp);
    G->div_hp = make_fastdiv(G->Hp);
    const long long total = static_cast<long long>(NB) * G->Hp * G->Wp;
    if (total >= (1ll << 30)) return false;
    G->total = static_cast<int>(total);
    G->num_tiles = (G->total + dtf::kTileM - 1) / dtf::kTileM;
    G->rows = dtf::kTileM + 2 * G->Wp + 2;
    if (G->rows > dtf::kHMaxRows) return false;
    // chunk stride: a multiple of 128 bytes.  (A stride of 128k + 16, which would spread the pair-of-chunk items
    // of one halo row over different banks, makes the tensor core's operand fetch 4x slower: measured 155 instead
    // of 41 cycles per MMA.)
    G->cs = ((G->rows * 16) + 127) & ~127;
    G->c8 = C / 8;
    return true;
}

static int prep_fwd(const DtfConvGemm* p, dtf::HaloGeom* Gout, int* smem_out);
static int prep_wgrad(const DtfConvWgrad* p, dtf::HaloGeom* Gout, int* smem_out, int* ny_out);

// returns 1 if launched, 0 if not eligible, <0 on error
extern "C" int dtf_conv3x3_try(const DtfConvGemm* p, cudaStream_t stream) {
    using namespace dtf;
    HaloGeom G{};
    int smem_bytes = 0;
    const int ok = prep_fwd(p, &G, &smem_bytes);
    if (ok != 1) return ok;
    static int attr_set = 0;
    if (!attr_set) {
        if (cudaFuncSetAttribute(conv3x3_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        attr_set = 1;
    }
    const int grid = G.num_tiles < halo_sms() ? G.num_tiles : halo_sms();
    if (G.c8 == 2) dtf_launch(conv3x3_kernel<2>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    else if (G.c8 == 4) dtf_launch(conv3x3_kernel<4>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    else dtf_launch(conv3x3_kernel<8>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    return cudaGetLastError() == cudaSuccess ? 1 : DTF_ERR_CUDA;
}

static int prep_fwd(const DtfConvGemm* p, dtf::HaloGeom* Gout, int* smem_out) {
    using namespace dtf;
    const DtfGather& a = p->a;
    if (halo_disabled()) return 0;
    if (a.KH != 3 || a.KW != 3 || a.stride != 1 || a.pad != 1) return 0;
    if (a.dtype != DTF_BF16) return 0;                           // fp32 tensors: generic kernels (conv_sm100.cu)
    if (a.C % 16 != 0 || a.C > 64 || p->N != a.C) return 0;      // kernel is specialised for Cout == Cin
    if (a.Hs != p->Ho || a.Ws != p->Wo) return 0;
    HaloGeom G{};
    if (!fill_geom(&G, p->NB, p->Ho, p->Wo, a.C)) return 0;
    G.sign = a.transposed ? -1 : 1;
    G.stage_bytes = (G.c8 * G.cs + 127) & ~127;
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    G.smem_b_off = (header + 1023) & ~1023;
    G.smem_a_off = (G.smem_b_off + p->kchunks * p->N * 16 + 1023) & ~1023;
    const int smem_bytes = G.smem_a_off + kHStages * G.stage_bytes + 256;
    if (smem_bytes > 227 * 1024) return 0;
    if (G.c8 != 2 && G.c8 != 4 && G.c8 != 8) return 0;
    *Gout = G;
    *smem_out = smem_bytes;
    return 1;
}

extern "C" int dtf_conv3x3_wgrad_try(const DtfConvWgrad* p, cudaStream_t stream) {
    using namespace dtf;
    HaloGeom G{};
    int smem_bytes = 0, ny = 1;
    const int ok = prep_wgrad(p, &G, &smem_bytes, &ny);
    if (ok != 1) return ok;
    static int attr_set = 0;
    if (!attr_set) {
        if (cudaFuncSetAttribute(conv3x3_wgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_wgrad_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_wgrad_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        attr_set = 1;
    }
    int gx = halo_sms() / ny;
    if (gx > G.num_tiles) gx = G.num_tiles;
    if (gx < 1) gx = 1;
    dim3 grid(gx, ny);
    if (G.c8 == 2) dtf_launch_ex(G.bulk != 0, conv3x3_wgrad_kernel<2>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    else if (G.c8 == 4) dtf_launch_ex(G.bulk != 0, conv3x3_wgrad_kernel<4>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    else dtf_launch_ex(G.bulk != 0, conv3x3_wgrad_kernel<8>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *p, G);
    return cudaGetLastError() == cudaSuccess ? 1 : DTF_ERR_CUDA;
}

static int prep_wgrad(const DtfConvWgrad* p, dtf::HaloGeom* Gout, int* smem_out, int* ny_out) {
    using namespace dtf;
    const DtfGather& x = p->x;
    if (halo_disabled()) return 0;
    if (x.KH != 3 || x.KW != 3 || x.stride != 1 || x.pad != 1) return 0;
    if (x.dtype != DTF_BF16) return 0;
    if (x.C % 16 != 0 || x.C > 64 || p->Cout % 8 != 0 || p->Cout > 128) return 0;
    if (x.Hs != p->Ho || x.Ws != p->Wo) return 0;
    HaloGeom G{};
    if (!fill_geom(&G, p->NB, p->Ho, p->Wo, x.C)) return 0;
    G.sign = 1;
    G.gy_c8 = p->Cout / 8;
    // TMEM holds ntaps * Cin fp32 columns (<= 512): split the nine taps over blockIdx.y when needed
    const int tmem_taps = 512 / x.C;
    const int ny = tmem_taps >= 9 ? 1 : 2;
    G.ntaps = ny == 1 ? 9 : 5;
    if (ny == 2 && (9 - G.ntaps) * x.C > 512) return 0;
    // stage = [gy tile][im2col: max taps per CTA * C8 chunks][halo tile]
    const int max_taps = G.ntaps > 9 - G.ntaps ? G.ntaps : 9 - G.ntaps;
    // DTF_WGRAD_BULK=1 (16 channels only): build the im2col operand with per-tap shared -> shared bulk copies that
    // keep the halo's chunk stride, instead of with the gather threads.  Off by default: on B200 the copy engine
    // moved the 47 KB of a tile in 1.25 us (~24 B/clk) and needs ~40 ns per copy issued, so the fused backward
    // launch of a 16-channel layer took 28.2 us against 25.9 us with thread copies (2 KB copies per (tap, chunk)
    // were slower still).
    static int bulk_env = -1;
    if (bulk_env < 0) { const char* e = getenv("DTF_WGRAD_BULK"); bulk_env = e ? atoi(e) : 0; }
    G.bulk = (bulk_env != 0 && G.c8 == 2) ? 1 : 0;
    G.tap0 = max_taps * G.c8 * (G.bulk ? G.cs : kChunkBytes);            // im2col area bytes
    G.stage_bytes = (G.gy_c8 * kChunkBytes + G.tap0 + G.c8 * G.cs + 127) & ~127;
    const int header = 1024 + static_cast<int>(sizeof(SmemConsts));
    G.smem_a_off = (header + 1023) & ~1023;
    // the gy operand descriptor spans 16 channel chunks (32 KB) from its base: the im2col area that follows
    // it inside the stage is at least that large whenever gy_c8 + chunks >= 16, else keep slack at the end
    const int slack = 16 * kChunkBytes;
    int stages = (227 * 1024 - G.smem_a_off - slack) / G.stage_bytes;
    if (stages > kHStages) stages = kHStages;
    if (stages < 1) return 0;
    G.sign = stages;                                                      // (re-used field: operand stages)
    const int smem_bytes = G.smem_a_off + stages * G.stage_bytes + slack;
    if (G.c8 != 2 && G.c8 != 4 && G.c8 != 8) return 0;
    if (G.gy_c8 != 2 && G.gy_c8 != 4 && G.gy_c8 != 8) return 0;
    *Gout = G;
    *smem_out = smem_bytes;
    *ny_out = ny;
    return 1;
}

// Fused backward launch (dgrad + wgrad of one 3x3 / stride-1 convolution).  Both parameter blocks must
// already be prepared by dtf_conv_gemm / dtf_conv_wgrad's front halves (kchunks etc.), which is what
// dtf_conv_bwd_fused in conv_sm100.cu does.  Returns 1 if launched, 0 if not eligible.
extern "C" int dtf_conv3x3_bwd_fused_try(const DtfConvGemm* pd, const DtfConvWgrad* pw, cudaStream_t stream) {
    using namespace dtf;
    HaloGeom Gd{}, Gw{};
    int smem_d = 0, smem_w = 0, ny = 1;
    int ok = prep_fwd(pd, &Gd, &smem_d);
    if (ok != 1) return ok;
    ok = prep_wgrad(pw, &Gw, &smem_w, &ny);
    if (ok != 1) return ok;
    if (Gd.c8 != Gw.c8) return 0;
    static int attr_set = 0;
    if (!attr_set) {
        if (cudaFuncSetAttribute(conv3x3_bwd_fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_bwd_fused_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(conv3x3_bwd_fused_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
            return DTF_ERR_CUDA;
        attr_set = 1;
    }
    const int sms = halo_sms();
    // Split the SMs between the two halves so that the slower half finishes as early as possible.  Both halves
    // walk whole tiles, so the cost of a split is max(ceil(T / n_d) * t_d, ceil(T / gx) * t_w); the per-tile
    // weights were fitted to a sweep of the split on B200 (bench/conv_micro.py with DTF_BWD_SPLIT=percent):
    // 16 ch: 1.7 / 1.28 us per tile, 32 ch: 3.0 / 3.0, 64 ch (two tap groups): 9.7 / 5.7 (re-fitted for the 24-warp
    // layout: finer gather items, two epilogue sets, alternating weight-gradient gather groups).  A fixed 45 % costs
    // 10-15 % on every stage.
    static int pct = -1;
    if (pct < 0) { const char* e = getenv("DTF_BWD_SPLIT"); pct = e ? atoi(e) : 0; if (pct < 5 || pct > 95) pct = 0; }
    const double t_d = Gd.c8 == 2 ? 1.7 : (Gd.c8 == 4 ? 3.0 : 9.7), t_w = Gd.c8 == 2 ? 1.28 : (Gd.c8 == 4 ? 3.0 : 5.7);
    int n_d = (sms * 45) / 100;
    if (pct > 0) {
        n_d = (sms * pct) / 100;
    } else {
        double best = 1e30, best2 = 1e30;
        for (int cand = 8; cand <= sms - 8 * ny; ++cand) {
            const int g = (sms - cand) / ny;
            const double cd = static_cast<double>((Gd.num_tiles + cand - 1) / cand) * t_d;
            const double cw = static_cast<double>((Gw.num_tiles + g - 1) / g) * t_w;
            const double c1 = cd > cw ? cd : cw, c2 = cd + cw;
            if (c1 < best - 1e-9 || (c1 < best + 1e-9 && c2 < best2 - 1e-9)) { best = c1; best2 = c2; n_d = cand; }
        }
    }
    if (n_d > Gd.num_tiles) n_d = Gd.num_tiles;
    if (n_d < 1) n_d = 1;
    int gx = (sms - n_d) / ny;
    if (gx > Gw.num_tiles) gx = Gw.num_tiles;
    if (gx < 1) gx = 1;
    const int grid = n_d + gx * ny;
    const int smem_bytes = smem_d > smem_w ? smem_d : smem_w;
    if (Gd.c8 == 2) dtf_launch_ex(Gw.bulk != 0, conv3x3_bwd_fused_kernel<2>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *pd, Gd, *pw, Gw, n_d, gx);
    else if (Gd.c8 == 4) dtf_launch_ex(Gw.bulk != 0, conv3x3_bwd_fused_kernel<4>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *pd, Gd, *pw, Gw, n_d, gx);
    else dtf_launch_ex(Gw.bulk != 0, conv3x3_bwd_fused_kernel<8>, dim3(grid), dim3(kHThreads), smem_bytes, stream, *pd, Gd, *pw, Gw, n_d, gx);
    return cudaGetLastError() == cudaSuccess ? 1 : DTF_ERR_CUDA;
}

#ifdef DTF_TIMING
// Copies the 64 timeline slots (globaltimer ns, written by CTA 0) to the host.
extern "C" int dtf_debug_timeline(unsigned long long* out) {
    return cudaMemcpyFromSymbol(out, dtf::g_timeline, sizeof(unsigned long long) * 128) == cudaSuccess ? 0 : -1;
}
#endif
