// Communication-bound hot paths as single kernels over NVLink peer memory:
//
//   exploit transfer   winner -> loser copy of the whole member state (weights + optimizer slots +
//                      BN moving statistics + step counter) read straight out of the winner GPU's
//                      memory with 16-byte peer loads, fused with writing the loser's new (explored)
//                      hyper-parameter block, resetting its per-round accumulators and publishing a
//                      completion epoch.   Replaces the master's `rm` + `cp` over a shared file system
//                      (reference pbt_cluster.py:145-147,168-181).
//   score board        per-member eval score pushed into the master rank's memory with a release
//                      store of the round epoch.  Replaces the pickled GET gather as the source of the
//                      ranking input (reference pbt_cluster.py:114-128).
//   all-reduce + update one-shot all-reduce of the gradient arena over peer loads fused with 1/W
//                      scaling, regulariser gradient, the optimizer rule and the bf16 weight re-pack
//                      (reference: dormant MirroredStrategy path, distribution_utils.py:41-47, followed
//                      by TF's per-variable Apply* ops, resnet_run_loop.py:552-586).
//
// Memory model: cross-GPU flags are written with st.release.sys after a __threadfence_system and
// polled with ld.acquire.sys; all waits are watchdog-protected (trap, never hang).
#include "common.cuh"
#include "dtf_b200.h"

#include <mutex>

namespace dtf {

DTF_DEVINL void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
DTF_DEVINL unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
DTF_DEVINL uint4 ld_peer16(const uint4* p) {
    uint4 r;   // plain (coherent) load: peer memory bypasses the local L2 anyway
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
DTF_DEVINL void wait_flag_ge(const unsigned long long* p, unsigned long long want, unsigned int site) {
    if (ld_acquire_sys(p) >= want) return;
    const uint64_t t0 = globaltimer_ns();
    while (ld_acquire_sys(p) < want) {
        if (globaltimer_ns() - t0 > kWatchdogNs) device_fail(0x2000u | site);
    }
}

// ---------------------------------------------------------------------------------------
// exploit: dst[0..bytes) = src[0..bytes) except [hp_off, hp_off+hp_bytes) := new hyper-parameters,
//          zero[0..zero_bytes) = 0, then *done_flag = epoch (release)
// `src` may be a peer-mapped pointer (pull) or `dst` may be (push): the code is the same.
// ---------------------------------------------------------------------------------------
constexpr int kCopyUnroll = 8;

__global__ void __launch_bounds__(512)
exploit_copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src, long long n_vec,
                    long long hp_vec_off, int hp_vecs, const __grid_constant__ DtfHparams hp,
                    uint4* __restrict__ zero, long long zero_vecs,
                    unsigned long long* done_flag, unsigned long long epoch, unsigned int* arrive_counter,
                    const unsigned long long* wait_flag, unsigned long long wait_value) {
    // device-side ordering against the producer of `src` (the winner's snapshot publish): acquire its epoch flag
    // before the first load.  Every CTA waits for itself, so no CTA depends on another one being resident.
    if (wait_flag != nullptr) {
        if (threadIdx.x == 0) wait_flag_ge(wait_flag, wait_value, 3);
        __syncthreads();
    }
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    long long i = tid;
    for (; i + (kCopyUnroll - 1) * nthreads < n_vec; i += kCopyUnroll * nthreads) {
        uint4 v[kCopyUnroll];
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u) v[u] = ld_peer16(src + i + u * nthreads);
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u) {
            const long long k = i + u * nthreads;
            if (k < hp_vec_off || k >= hp_vec_off + hp_vecs) dst[k] = v[u];
        }
    }
    for (; i < n_vec; i += nthreads)
        if (i < hp_vec_off || i >= hp_vec_off + hp_vecs) dst[i] = ld_peer16(src + i);
    // explore write: the (already perturbed) hyper-parameter block replaces the winner's
    if (hp_vecs > 0 && tid < hp_vecs) {
        const uint4* hv = reinterpret_cast<const uint4*>(&hp);
        constexpr int kHpVecs = sizeof(DtfHparams) / 16;
        dst[hp_vec_off + tid] = tid < kHpVecs ? hv[tid] : make_uint4(0, 0, 0, 0);
    }
    // loser state reset (per-round metric accumulators)
    for (long long z = tid; z < zero_vecs; z += nthreads) zero[z] = make_uint4(0, 0, 0, 0);
    // publish completion: last CTA to finish releases the flag
    if (done_flag != nullptr) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned int prev = atomicAdd(arrive_counter, 1u);
            if (prev == gridDim.x - 1) {
                *arrive_counter = 0;
                __threadfence_system();
                st_release_sys(done_flag, epoch);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// score board: slot = {float score; uint32 count; uint64 epoch}
// ---------------------------------------------------------------------------------------
__global__ void score_publish_kernel(DtfScoreSlot* __restrict__ board_slot, const int* __restrict__ correct,
                                     int total, unsigned long long epoch, int reset_counter, int* correct_rw) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const int c = *correct;
        board_slot->hits = static_cast<unsigned int>(c);
        board_slot->seen = static_cast<unsigned int>(total);
        __threadfence_system();
        st_release_sys(&board_slot->epoch, epoch);
        if (reset_counter) *correct_rw = 0;
    }
}

// start of a member's evaluation: the slot's counters restart at zero (the epoch keeps its last value)
__global__ void score_reset_kernel(DtfScoreSlot* __restrict__ board_slot) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        board_slot->hits = 0u;
        board_slot->seen = 0u;
    }
}

// end of a member's evaluation: every evaluation kernel of the round has completed (stream order), their
// system-scope reductions into the slot are performed; publish the epoch.  A non-finite loss turns the entry
// into the "dropped" marker so that the master does not wait for a member its worker is about to remove.
__global__ void score_release_kernel(DtfScoreSlot* __restrict__ board_slot, unsigned long long epoch,
                                     const float* __restrict__ loss, int dropped) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        bool bad = dropped != 0;
        if (loss != nullptr) {
            const float l = *loss;
            bad = bad || !(l == l) || fabsf(l) > 3.0e38f;
        }
        if (bad) board_slot->hits = DTF_SCORE_DROPPED;
        __threadfence_system();
        st_release_sys(&board_slot->epoch, epoch);
    }
}

// ---------------------------------------------------------------------------------------
// fused one-shot all-reduce + optimizer update (data-parallel member, W = 2 or 4 GPUs)
// ---------------------------------------------------------------------------------------
DTF_DEVINL float schedule_lr2(const DtfHparams& hp, long long step) {
    int i = 0;
    while (i < hp.n_boundaries && step > hp.boundaries[i]) ++i;
    return hp.lr_values[i];
}

DTF_DEVINL void pack_store2(const DtfOptSegment& s, dtf_bf16* pack, int local, float w) {
    if (s.pack == 2) { pack[s.fwd_pack_off + local] = __float2bfloat16(w); return; }
    const int ci = local % s.Cin;
    const int rest = local / s.Cin;
    const int tap = rest % s.taps;
    const int co = rest / s.taps;
    if (s.pack == 3) {            // fp32 operands (tf32 MMA): [K/4][N][4] floats (elementwise.cu: pack_store)
        float* pf = reinterpret_cast<float*>(pack);
        if (s.fwd_pack_off >= 0) {
            const int k = tap * s.Cin + ci;
            pf[s.fwd_pack_off + ((k >> 2) * s.Cout + co) * 4 + (k & 3)] = w;
        }
        if (s.dgrad_pack_off >= 0) {
            const int k = tap * s.Cout + co;
            pf[s.dgrad_pack_off + ((k >> 2) * s.Cin + ci) * 4 + (k & 3)] = w;
        }
        return;
    }
    const dtf_bf16 h = __float2bfloat16(w);
    if (s.fwd_pack_off >= 0) {
        const int k = tap * s.Cin + ci;
        pack[s.fwd_pack_off + ((k >> 3) * s.Cout + co) * 8 + (k & 7)] = h;
    }
    if (s.dgrad_pack_off >= 0) {
        const int k = tap * s.Cout + co;
        pack[s.dgrad_pack_off + ((k >> 3) * s.Cin + ci) * 8 + (k & 7)] = h;
    }
}

__global__ void __launch_bounds__(256)
allreduce_update_kernel(const __grid_constant__ DtfAllreduceUpdate P) {
    __shared__ DtfHparams hp;
    const int W = P.world;
    // The epoch of this exchange is a device-resident counter (bumped by dtf_dp_step_begin), so a captured
    // CUDA graph of the step replays with the right epoch; P.epoch is only an offset.  It is NOT the step
    // counter: that one is part of the member state and jumps when the member adopts a winner or restores.
    const long long step = *P.step;
    const unsigned long long epoch = *P.epoch_ctr + P.epoch;
    // ---- entry barrier: tell every peer that my gradients for `epoch` are complete, then wait for theirs
    if (blockIdx.x == 0 && threadIdx.x < W) {
        __threadfence_system();
        st_release_sys(P.peer_ready[threadIdx.x] + P.rank, epoch);     // my slot in peer's ready[]
    }
    uint64_t t_begin = 0, t_ready = 0, t_loaded = 0;
    if (P.timing != nullptr && blockIdx.x == 0 && threadIdx.x == 0) t_begin = globaltimer_ns();
    if (threadIdx.x == 0) hp = *P.hp;
    if (threadIdx.x < W) wait_flag_ge(P.local_ready + threadIdx.x, epoch, 1);
    __syncthreads();
    if (P.timing != nullptr && blockIdx.x == 0 && threadIdx.x == 0) t_ready = globaltimer_ns();
    const float lr = schedule_lr2(hp, step - 1);
    float lr_adam = lr;
    if (hp.optimizer == DTF_OPT_ADAM) {
        const float t = static_cast<float>(step);
        lr_adam = lr * sqrtf(1.f - powf(0.999f, t)) / (1.f - powf(0.9f, t));
    }
    const int2 ch = reinterpret_cast<const int2*>(P.chunks)[blockIdx.x];
    const DtfOptSegment s = P.segs[ch.x];
    const float gs = P.grad_scale * hp.inv_loss_scale;
    // One-shot: every rank reads all replicas.  A thread owns four consecutive elements so that the remote reads
    // are 16-byte loads over NVLink (a chunk is 1024 elements = 256 threads x 4); segment offsets are multiples of
    // four elements, only a segment's last few elements fall back to scalar loads.
    const int local0 = ch.y + 4 * static_cast<int>(threadIdx.x);
    const bool vec = ((s.offset | ch.y) & 3) == 0 && local0 + 3 < s.count;
    float gsum[4] = {0.f, 0.f, 0.f, 0.f};
    if (vec) {
#pragma unroll 4
        for (int w = 0; w < W; ++w) {
            const float4 r = *reinterpret_cast<const float4*>(P.peer_grads[w] + s.offset + local0);
            gsum[0] += r.x; gsum[1] += r.y; gsum[2] += r.z; gsum[3] += r.w;
        }
    } else {
        for (int e = 0; e < 4; ++e)
            if (local0 + e < s.count)
                for (int w = 0; w < W; ++w) gsum[e] += P.peer_grads[w][s.offset + local0 + e];
    }
    if (P.timing != nullptr && blockIdx.x == 0 && threadIdx.x == 0) t_loaded = globaltimer_ns();
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int local = local0 + e;
        if (local >= s.count) break;
        const int idx = s.offset + local;
        float g = gsum[e] * gs;
        const float w0 = P.params[idx];
        if (s.regularized) {
            if (hp.regularizer == DTF_REG_L1 || hp.regularizer == DTF_REG_L1_L2)
                g += hp.weight_decay * (w0 > 0.f ? 1.f : (w0 < 0.f ? -1.f : 0.f));
            if (hp.regularizer == DTF_REG_L2 || hp.regularizer == DTF_REG_L1_L2) g += hp.weight_decay * w0;
        }
        float s0 = P.slot0[idx], s1 = P.slot1[idx], nw;
        switch (hp.optimizer) {
            case DTF_OPT_GD: nw = w0 - lr * g; break;
            case DTF_OPT_MOMENTUM: s0 = hp.momentum * s0 + g; nw = w0 - lr * s0; break;
            case DTF_OPT_ADAGRAD: s0 += g * g; nw = w0 - lr * g * rsqrtf(s0); break;
            case DTF_OPT_ADADELTA: {
                s0 = 0.95f * s0 + 0.05f * g * g;
                const float upd = sqrtf(s1 + 1e-8f) * rsqrtf(s0 + 1e-8f) * g;
                s1 = 0.95f * s1 + 0.05f * upd * upd;
                nw = w0 - lr * upd;
                break;
            }
            case DTF_OPT_ADAM:
                s0 = 0.9f * s0 + 0.1f * g;
                s1 = 0.999f * s1 + 0.001f * g * g;
                nw = w0 - lr_adam * s0 / (sqrtf(s1) + 1e-8f);
                break;
            default:
                s0 = hp.grad_decay * s0 + (1.f - hp.grad_decay) * g * g;
                s1 = hp.momentum * s1 + lr * g * rsqrtf(s0 + 1e-10f);
                nw = w0 - s1;
        }
        P.params[idx] = nw;
        P.slot0[idx] = s0;
        P.slot1[idx] = s1;
        if (s.pack) pack_store2(s, P.pack, local, nw);
    }
    // device-timestamped accounting of the exchange (CTA 0): time blocked on the peers' ready flags (= exposed
    // communication: it is zero when every replica's backward finished together), time of the remote loads, and the
    // whole kernel -- accumulated in nanoseconds, read back by parallel/dp.py
    if (P.timing != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
        const uint64_t t_end = globaltimer_ns();
        atomicAdd(P.timing + 0, static_cast<unsigned long long>(t_ready - t_begin));
        atomicAdd(P.timing + 1, static_cast<unsigned long long>(t_loaded - t_ready));
        atomicAdd(P.timing + 2, static_cast<unsigned long long>(t_end - t_begin));
        atomicAdd(P.timing + 3, 1ull);
    }
    // ---- exit signal: once every CTA of this rank has consumed the peers' gradients, tell the peers
    // that their gradient arenas may be zeroed for the next step (checked by dtf_dp_step_begin)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int prev = atomicAdd(P.arrive_counter, 1u);
        if (prev == gridDim.x - 1) {
            *P.arrive_counter = 0;
            __threadfence_system();
            for (int w = 0; w < W; ++w) st_release_sys(P.peer_done[w] + P.rank, epoch);
        }
    }
}

// step begin for a DP member: wait until every peer has finished reading my gradients of the
// previous step, then zero the accumulators and bump the step counter
// (single CTA grid-stride would be slow; instead every CTA waits, and the LAST CTA to finish bumps the
// counter so that no CTA can observe the bumped value while computing its wait target)
__global__ void __launch_bounds__(256)
dp_step_begin_kernel(float4* __restrict__ zero, long long n_vec, long long* __restrict__ step_ptr,
                     unsigned long long* __restrict__ epoch_ctr,
                     const unsigned long long* __restrict__ local_done, int world, unsigned long long epoch_offset,
                     unsigned int* __restrict__ arrive_counter) {
    // peers must have consumed the gradients of the previous exchange (= current counter value)
    const unsigned long long prev_epoch = *epoch_ctr + epoch_offset;
    if (threadIdx.x < world) wait_flag_ge(local_done + threadIdx.x, prev_epoch, 2);
    __syncthreads();
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec;
         i += (long long)gridDim.x * blockDim.x)
        zero[i] = z;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int prev = atomicAdd(arrive_counter, 1u);
        if (prev == gridDim.x - 1) {
            *arrive_counter = 0;
            *step_ptr += 1;
            *epoch_ctr += 1;
        }
    }
}

}  // namespace dtf

#define DTF_CHECK_LAUNCH() (cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA)

extern "C" int dtf_exploit_copy_ordered(void* dst, const void* src, long long bytes, long long hp_offset,
                                        const DtfHparams* new_hp, void* zero_region, long long zero_bytes,
                                        unsigned long long* done_flag, unsigned long long epoch,
                                        unsigned int* arrive_counter, const unsigned long long* wait_flag,
                                        unsigned long long wait_value, int max_ctas, cudaStream_t stream) {
    if ((bytes & 15) || (zero_bytes & 15) || (hp_offset & 15)) return DTF_ERR_ARG;
    if (bytes < 0 || zero_bytes < 0) return DTF_ERR_ARG;
    if (((uintptr_t)dst & 15) || ((uintptr_t)src & 15)) return DTF_ERR_ARG;
    if (done_flag != nullptr && arrive_counter == nullptr) return DTF_ERR_ARG;
    // the hyper-parameter block must lie inside the copied range, or the kernel would write past `dst`
    if (new_hp != nullptr && (hp_offset < 0 || hp_offset + (long long)((sizeof(DtfHparams) + 15) / 16 * 16) > bytes))
        return DTF_ERR_ARG;
    const long long n_vec = bytes / 16;
    const int hp_vecs = new_hp ? static_cast<int>((sizeof(DtfHparams) + 15) / 16) : 0;
    DtfHparams hp_val;
    if (new_hp) hp_val = *new_hp; else memset(&hp_val, 0, sizeof(hp_val));
    long long want = (n_vec + 512LL * dtf::kCopyUnroll - 1) / (512LL * dtf::kCopyUnroll);
    if (want < 1) want = 1;
    int cap = max_ctas > 0 ? max_ctas : 148 * 2;
    const int grid = static_cast<int>(want < cap ? want : cap);
    dtf::exploit_copy_kernel<<<grid, 512, 0, stream>>>(
        reinterpret_cast<uint4*>(dst), reinterpret_cast<const uint4*>(src), n_vec,
        new_hp ? hp_offset / 16 : -1, hp_vecs, hp_val, reinterpret_cast<uint4*>(zero_region), zero_bytes / 16,
        done_flag, epoch, arrive_counter, wait_flag, wait_value);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_exploit_copy(void* dst, const void* src, long long bytes, long long hp_offset,
                                const DtfHparams* new_hp, void* zero_region, long long zero_bytes,
                                unsigned long long* done_flag, unsigned long long epoch,
                                unsigned int* arrive_counter, int max_ctas, cudaStream_t stream) {
    return dtf_exploit_copy_ordered(dst, src, bytes, hp_offset, new_hp, zero_region, zero_bytes, done_flag, epoch,
                                    arrive_counter, nullptr, 0ull, max_ctas, stream);
}

extern "C" int dtf_score_reset(DtfScoreSlot* board_slot, cudaStream_t stream) {
    if (board_slot == nullptr) return DTF_ERR_ARG;
    dtf::score_reset_kernel<<<1, 32, 0, stream>>>(board_slot);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_score_release(DtfScoreSlot* board_slot, unsigned long long epoch, const float* loss, int dropped,
                                 cudaStream_t stream) {
    if (board_slot == nullptr) return DTF_ERR_ARG;
    dtf::score_release_kernel<<<1, 32, 0, stream>>>(board_slot, epoch, loss, dropped);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_score_publish(DtfScoreSlot* board_slot, const int* correct, int total,
                                 unsigned long long epoch, int reset_counter, int* correct_rw,
                                 cudaStream_t stream) {
    dtf::score_publish_kernel<<<1, 32, 0, stream>>>(board_slot, correct, total, epoch, reset_counter, correct_rw);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_allreduce_update(const DtfAllreduceUpdate* p, cudaStream_t stream) {
    if (p->world < 1 || p->world > DTF_MAX_DP || p->epoch_ctr == nullptr) return DTF_ERR_ARG;
    dtf::allreduce_update_kernel<<<p->n_chunks, 256, 0, stream>>>(*p);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_dp_step_begin(float* zero_base, long long zero_floats, long long* step,
                                 unsigned long long* epoch_ctr,
                                 const unsigned long long* local_done, int world, unsigned long long epoch_offset,
                                 unsigned int* arrive_counter, cudaStream_t stream) {
    if (zero_floats % 4 != 0 || ((uintptr_t)zero_base & 15) != 0 || epoch_ctr == nullptr) return DTF_ERR_ARG;
    const long long n_vec = zero_floats / 4;
    long long blocks = (n_vec + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    dtf::dp_step_begin_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(
        reinterpret_cast<float4*>(zero_base), n_vec, step, epoch_ctr, local_done, world, epoch_offset, arrive_counter);
    return DTF_CHECK_LAUNCH();
}

// ---------------------------------------------------------------------------------------
// symmetric memory plumbing (host): allocation + CUDA IPC handles.  These DO allocate.
// ---------------------------------------------------------------------------------------
extern "C" int dtf_symm_alloc(long long bytes, void** out) {
    void* p = nullptr;
    if (cudaMalloc(&p, static_cast<size_t>(bytes)) != cudaSuccess) return DTF_ERR_CUDA;
    if (cudaMemset(p, 0, static_cast<size_t>(bytes)) != cudaSuccess) return DTF_ERR_CUDA;
    *out = p;
    return DTF_OK;
}
extern "C" int dtf_symm_free(void* p) { return cudaFree(p) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA; }
extern "C" int dtf_ipc_get_handle(void* p, unsigned char* handle64) {
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) return DTF_ERR_CUDA;
    memcpy(handle64, &h, sizeof(h));
    return DTF_OK;
}
extern "C" int dtf_ipc_open_handle(const unsigned char* handle64, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) return DTF_ERR_CUDA;
    *out = p;
    return DTF_OK;
}
extern "C" int dtf_ipc_close(void* p) { return cudaIpcCloseMemHandle(p) == cudaSuccess ? DTF_OK : DTF_ERR_CUDA; }
extern "C" int dtf_enable_peer(int peer_device) {
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return DTF_OK; }
    return e == cudaSuccess ? DTF_OK : DTF_ERR_CUDA;
}
extern "C" int dtf_last_cuda_error(char* buf, int n) {
    const char* s = cudaGetErrorString(cudaPeekAtLastError());
    snprintf(buf, n, "%s", s);
    return DTF_OK;
}

// =========================================================================================
// dtf_exploit_copy_noalloc — the exploit transfer for callers that must not allocate.
//
// Who needs it: a worker adopts a winner's state while the other members that share its GPU
// (or its process) are in the middle of a training round.  Any device allocation at that point
// is a problem — cudaMalloc / cudaFree synchronise the whole device and would stall every
// member's stream, a caller that embeds this library may have handed all device memory to its
// own pool, and a caller that sits inside a fixed memory budget simply has nowhere to put the
// kernel's bookkeeping.  The regular entry point makes such a caller bring 4 bytes of zeroed
// device scratch (the CTA arrival counter behind the completion flag); this variant owns that
// scratch itself.
//
// The scratch is the file-scope array below: kNoAllocScratchBytes (256) bytes, present once in
// every GPU context that loads the library and held until the process exits.  Only the first
// 4 bytes are used today (the arrival counter); the rest is reserved.  Nothing else changes:
// inputs and outputs are still the caller's.
//
// Rules that follow from sharing one array:
//   * calls must not overlap on the same GPU (two copies in flight on different streams would
//     count arrivals in the same cell and could publish a completion epoch early or never);
//     serialise them on one stream or with events;
//   * the counter is cleared by a memset on `stream` at the start of every call, so a previous
//     call that was aborted cannot poison the next one;
//   * the first call on each device resolves the array's device address (with that device
//     current) and caches it — make that first call outside any stream capture;
//   * `bytes` above DTF_NOALLOC_MAX_BYTES is refused with DTF_ERR_LIMIT (the message is available
//     from dtf_noalloc_last_error); larger states go through dtf_exploit_copy.
// No cudaMalloc / cudaFree / allocating library call is made here.
// =========================================================================================
namespace dtf {
constexpr size_t kNoAllocScratchBytes = 256;
static __device__ __align__(256) unsigned char g_exploit_noalloc_scratch[kNoAllocScratchBytes];
}  // namespace dtf

static const int kNoAllocMaxDevices = 64;
static void* g_noalloc_scratch_ptr[kNoAllocMaxDevices];      // device address of the array, per device
// Several member threads of one worker may call in at once (worker.members_in_flight): the pointer cache is
// guarded by a mutex and the error text is kept per calling thread, so neither can be torn.
static std::mutex g_noalloc_mutex;
static thread_local char g_noalloc_error[256] = "";

extern "C" const char* dtf_noalloc_last_error(void) { return g_noalloc_error; }

// In host code the array's name is only a stand-in for the symbol; the address a kernel may
// dereference has to be asked for, once per device, with that device current.
static int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error), "dtf_exploit_copy_noalloc: no usable current device (%d)", dev);
        return DTF_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, dtf::g_exploit_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            snprintf(g_noalloc_error, sizeof(g_noalloc_error),
                     "dtf_exploit_copy_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return DTF_ERR_CUDA;
        }
        g_noalloc_scratch_ptr[dev] = p;
    }
    *out = g_noalloc_scratch_ptr[dev];
    return DTF_OK;
}

extern "C" int dtf_exploit_copy_noalloc(void* dst, const void* src, long long bytes, long long hp_offset,
                                        const DtfHparams* new_hp, void* zero_region, long long zero_bytes,
                                        unsigned long long* done_flag, unsigned long long epoch,
                                        cudaStream_t stream) {
    if (bytes < 0 || bytes > DTF_NOALLOC_MAX_BYTES) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error),
                 "dtf_exploit_copy_noalloc: %lld bytes exceeds the supported maximum of %lld bytes; "
                 "use dtf_exploit_copy for larger states", bytes, (long long)DTF_NOALLOC_MAX_BYTES);
        return DTF_ERR_LIMIT;
    }
    if ((bytes & 15) || (zero_bytes & 15) || (hp_offset & 15) || ((uintptr_t)dst & 15) || ((uintptr_t)src & 15)) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error),
                 "dtf_exploit_copy_noalloc: pointers, sizes and the hyper-parameter offset must be multiples of 16 bytes");
        return DTF_ERR_ARG;
    }
    if (zero_bytes < 0 ||
        (new_hp != nullptr && (hp_offset < 0 || hp_offset + (long long)((sizeof(DtfHparams) + 15) / 16 * 16) > bytes))) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error),
                 "dtf_exploit_copy_noalloc: the hyper-parameter block [%lld, +%zu) must lie inside the %lld copied bytes "
                 "and zero_bytes must not be negative", hp_offset, sizeof(DtfHparams), bytes);
        return DTF_ERR_ARG;
    }
    void* scratch = nullptr;
    const int rc = noalloc_scratch_for_current_device(&scratch);
    if (rc != DTF_OK) return rc;
    // the arrival counter has to start at zero; whatever an earlier call left behind is wiped here
    if (cudaMemsetAsync(scratch, 0, sizeof(unsigned int), stream) != cudaSuccess) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error), "dtf_exploit_copy_noalloc: clearing the scratch failed: %s",
                 cudaGetErrorString(cudaGetLastError()));
        return DTF_ERR_CUDA;
    }
    const long long n_vec = bytes / 16;
    const int hp_vecs = new_hp ? static_cast<int>((sizeof(DtfHparams) + 15) / 16) : 0;
    DtfHparams hp_val;
    if (new_hp) hp_val = *new_hp; else memset(&hp_val, 0, sizeof(hp_val));
    long long want = (n_vec + 512LL * dtf::kCopyUnroll - 1) / (512LL * dtf::kCopyUnroll);
    if (want < 1) want = 1;
    const int grid = static_cast<int>(want < 148 * 2 ? want : 148 * 2);
    dtf::exploit_copy_kernel<<<grid, 512, 0, stream>>>(
        reinterpret_cast<uint4*>(dst), reinterpret_cast<const uint4*>(src), n_vec,
        new_hp ? hp_offset / 16 : -1, hp_vecs, hp_val, reinterpret_cast<uint4*>(zero_region), zero_bytes / 16,
        done_flag, epoch, static_cast<unsigned int*>(scratch), nullptr, 0ull);
    if (cudaGetLastError() != cudaSuccess) {
        snprintf(g_noalloc_error, sizeof(g_noalloc_error), "dtf_exploit_copy_noalloc: kernel launch failed");
        return DTF_ERR_CUDA;
    }
    return DTF_OK;
}
