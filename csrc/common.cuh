// Shared device helpers for the sm_100a kernels: mbarrier, tcgen05 (TMEM alloc / MMA /
// commit / ld), proxy fences, UMMA descriptors, bf16 packing, watchdog-protected waits.
//
// Everything here is inline PTX for sm_100a; there is no fallback path.  The UMMA shared
// memory layout used throughout is the *no-swizzle interleaved* canonical layout
//      element (row r, k) of a tile  ->  chunk (k / 8) , row r , sub-element (k % 8)
//      byte address = chunk * chunk_stride + r * 16 + (k % 8) * 2
// i.e. "[k/8][row][8]" with 16-byte rows.  Viewed K-major (rows = M/N index) the
// descriptor is LBO = chunk_stride, SBO = 128; viewed MN-major (the 16-byte row is 8
// consecutive M/N elements, rows = K index) it is SBO = chunk_stride, LBO = 128.  The same
// gathered tile therefore serves the forward/dgrad GEMMs (K-major) and the wgrad GEMM
// (MN-major) without any transpose.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DTF_DEVINL __device__ __forceinline__

namespace dtf {

// ---------------------------------------------------------------------------------------
// error reporting from device code: print the failing site and trap (never hang the GPU)
// ---------------------------------------------------------------------------------------
// (never inlined: the printf set-up would otherwise be copied into every wait site -- tens of kilobytes of cold
// code inside the kernels' loops, and instruction-cache misses are a first-order cost for 10 us kernels)
static __device__ __noinline__ void device_fail(unsigned int code) {
    printf("[dtf_b200] device watchdog/assert: site 0x%x block (%d,%d) thread %d\n", code,
           (int)blockIdx.x, (int)blockIdx.y, (int)threadIdx.x);
    __trap();
}

DTF_DEVINL uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

constexpr uint64_t kWatchdogNs = 4000000000ull;   // 4 s: a wait this long is a protocol bug

// ---------------------------------------------------------------------------------------
// shared-memory addresses
// ---------------------------------------------------------------------------------------
DTF_DEVINL uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------
DTF_DEVINL void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
DTF_DEVINL void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
DTF_DEVINL void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
DTF_DEVINL void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
DTF_DEVINL bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Wait for the phase with the given parity to complete.  Traps (instead of hanging the GPU)
// if the barrier has not flipped within the watchdog window.
static __device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity, unsigned int site) {
    const uint64_t t0 = globaltimer_ns();
    while (!mbar_try_wait(bar, parity)) {
        if (globaltimer_ns() - t0 > kWatchdogNs) device_fail(0x1000u | site);
    }
}
DTF_DEVINL void mbar_wait(uint32_t bar, uint32_t parity, unsigned int site) {
    // a few polls inline (try_wait itself suspends the thread for a while), then the out-of-line watchdog loop
#pragma unroll 1
    for (int i = 0; i < 8; ++i)
        if (mbar_try_wait(bar, parity)) return;
    mbar_wait_slow(bar, parity, site);
}

// ---------------------------------------------------------------------------------------
// proxy / tcgen05 fences
// ---------------------------------------------------------------------------------------
// generic-proxy st.shared  ->  async-proxy readers (tcgen05.mma, bulk copies)
DTF_DEVINL void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
DTF_DEVINL void tc_fence_before_sync() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
DTF_DEVINL void tc_fence_after_sync() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// ---------------------------------------------------------------------------------------
// bulk (TMA) copy global -> shared, 1-D, completion on an mbarrier (SASS: UBLKCP)
// ---------------------------------------------------------------------------------------
DTF_DEVINL void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// bulk copy shared -> shared inside the CTA (the async copy engine moves the bytes, no thread touches them);
// completion is counted on an mbarrier like a global -> shared bulk copy.  16-byte aligned addresses and size.
DTF_DEVINL void bulk_s2s(uint32_t dst_smem, uint32_t src_smem, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst_smem), "r"(src_smem), "r"(bytes), "r"(bar) : "memory");
}

// ---------------------------------------------------------------------------------------
// TMEM allocation (one full warp executes these)
// ---------------------------------------------------------------------------------------
DTF_DEVINL void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
DTF_DEVINL void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__host__ __device__ constexpr uint32_t tmem_cols_pow2(uint32_t n) {
    return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : n <= 256 ? 256 : 512;
}

// ---------------------------------------------------------------------------------------
// UMMA descriptors
// ---------------------------------------------------------------------------------------
// Shared-memory matrix descriptor, SWIZZLE_NONE, sm_100 version bits set.
//   bits [0,14)  start address >> 4      bits [16,30) leading byte offset >> 4
//   bits [32,46) stride byte offset >> 4 bits [46,48) version = 1      bits [61,64) layout = 0
DTF_DEVINL uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= 1ull << 46;
    return d;
}
// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32.
//   [4,6) c_format (1 = F32)  [7,10) a_format (1 = BF16)  [10,13) b_format (1 = BF16)
//   [15] a_major (0 = K, 1 = MN)  [16] b_major  [17,23) N >> 3   [24,29) M >> 4
__host__ __device__ constexpr uint32_t umma_idesc_bf16(uint32_t m, uint32_t n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((n >> 3) << 17) | ((m >> 4) << 24);
}
// Instruction descriptor for kind::tf32 (fp32 storage, tf32 multiply, fp32 accumulate): a/b format 2.  One MMA
// covers K = 8 elements = the same 32 bytes per row as 16 bf16, so the operand byte layouts are shared.
__host__ __device__ constexpr uint32_t umma_idesc_tf32(uint32_t m, uint32_t n, bool a_mn, bool b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((n >> 3) << 17) | ((m >> 4) << 24);
}
DTF_DEVINL void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]   — issued by ONE thread.
DTF_DEVINL void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// Same with the descriptors given as (low word, shared high word) pairs and the accumulate flag as a compile-time
// constant: the issuing thread spends ~40 ns per MMA on operand set-up otherwise (64-bit adds, a setp per issue),
// twice what the tensor core needs for the small-N MMAs of the 16-channel layers.
template <bool kAccumulate>
DTF_DEVINL void umma_bf16_lohi(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc) {
    if (kAccumulate) {
        asm volatile(
            "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
            "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.eq.u32 p, 1, 1;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
            ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc) : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
            "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.ne.u32 p, 1, 1;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
            ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc) : "memory");
    }
}
// Make an mbarrier track completion of all previously issued MMAs of this thread
// (implies tcgen05.fence::before_thread_sync).
DTF_DEVINL void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// ---------------------------------------------------------------------------------------
// TMEM -> registers.  A warp reads its own 32-lane quadrant (lane = 32 * (warp % 4) + laneid).
// ---------------------------------------------------------------------------------------
DTF_DEVINL void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
DTF_DEVINL void tmem_ld8(uint32_t taddr, float* v) {
    uint32_t r[8];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// ---------------------------------------------------------------------------------------
// bf16 <-> fp32 packing on 16-byte vectors
// ---------------------------------------------------------------------------------------
DTF_DEVINL void unpack8(const uint4& q, float* f) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = __uint_as_float(w[i] << 16);
        f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
}
DTF_DEVINL uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
DTF_DEVINL uint4 pack8(const float* f) {
    return make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
}
DTF_DEVINL uint4 ldg16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
DTF_DEVINL void sts16(uint32_t addr, const uint4& v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// one elected lane of a fully converged warp (same lane every time for a full mask)
DTF_DEVINL bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---- programmatic dependent launch ------------------------------------------------------
// Every kernel of the training step is launched with the programmatic-stream-serialization
// attribute (dtf_launch below): its CTAs may be scheduled while the previous kernel is still
// draining, run their private set-up (barriers, TMEM allocation, shared tables), and block in
// griddep_wait() until the previous grid has completed and its writes are visible.  Rules kept by
// all kernels: nothing in global memory is read or written before griddep_wait(); every kernel
// executes griddep_wait() (completion is then transitive along the stream).
DTF_DEVINL void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
DTF_DEVINL void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// named barrier among a subset of warps
DTF_DEVINL void named_bar_sync(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace dtf

// Host side: launch with the programmatic-dependent-launch attribute (DTF_NO_PDL=1 turns it off).
#include <stdlib.h>
inline bool dtf_pdl_enabled() {
    static int on = -1;
    if (on < 0) on = getenv("DTF_NO_PDL") != nullptr ? 0 : 1;
    return on != 0;
}
template <typename... KArgs, typename... Args>
inline cudaError_t dtf_launch_ex(bool cluster_of_one, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (cluster_of_one) {      // explicit 1x1x1 cluster: makes the shared::cluster forms (shared -> shared bulk copy) legal
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = 1;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (dtf_pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t dtf_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    return dtf_launch_ex(false, kernel, grid, block, smem, stream, static_cast<Args&&>(args)...);
}
