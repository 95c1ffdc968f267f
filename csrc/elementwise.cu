// Non-GEMM kernels of the training step: BN-backward apply (+ residual gradient), the
// classifier head (final BN + ReLU + global average pool + dense + softmax cross-entropy,
// forward and backward in one launch), weight packing, the multi-tensor optimizer with the six
// TF-1.x update rules, step bookkeeping, and input preprocessing.
//
// Reference call sites replaced: K3/K4 BN backward, K7 reduce_mean, K8 dense,
// K9 softmax/argmax/CE/accuracy (resnet_model.py:537-552, resnet_run_loop.py:230-246,303),
// K10 regularisers (resnet_model.py:111-121), K11 optimizers (resnet_run_loop.py:552-586),
// K12 loss scaling, K13 LR schedule (resnet_run_loop.py:135-173), K14 preprocessing
// (cifar10_main.py:71-109).
#include "common.cuh"
#include "dtf_b200.h"

namespace dtf {

DTF_DEVINL void bn_consts(const DtfBnRef& bn, int c, float& scale, float& shift, float& mean, float& invstd) {
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

// ---------------------------------------------------------------------------------------
// g[p][c] = p0[c]*t + p1[c]*x + p2[c] (+ res)      BN backward, given (sum t, sum t*xhat)
// ---------------------------------------------------------------------------------------
// E = elements per 16-byte vector: 8 (bf16 tensors) or 4 (fp32 tensors)
template <int E>
DTF_DEVINL void unpack_vec(const uint4& q, float* f) {
    if (E == 8) unpack8(q, f);
    else { f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w); }
}
template <int E>
DTF_DEVINL uint4 pack_vec(const float* f) {
    if (E == 8) return pack8(f);
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}

template <int E>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const dtf_bf16* __restrict__ t, const dtf_bf16* __restrict__ x,
                    const dtf_bf16* __restrict__ res, dtf_bf16* __restrict__ out,
                    DtfBnRef bn, const float* __restrict__ bwd_stats, long long total_vec, int C) {
    griddep_launch_dependents();
    griddep_wait();
    __shared__ float p0[256], p1[256], p2[256];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float scale, shift, mean, invstd;
        bn_consts(bn, c, scale, shift, mean, invstd);
        const float mt = bwd_stats[c] * bn.inv_count;
        const float mtx = bwd_stats[C + c] * bn.inv_count;
        p0[c] = scale;
        p1[c] = -scale * invstd * mtx;
        p2[c] = scale * (mean * invstd * mtx - mt);
    }
    __syncthreads();
    const int cvec = C / E;
    const uint4* tv = reinterpret_cast<const uint4*>(t);
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    const uint4* rv = reinterpret_cast<const uint4*>(res);
    uint4* ov = reinterpret_cast<uint4*>(out);
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total_vec;
         v += (long long)gridDim.x * blockDim.x) {
        const int cb = static_cast<int>(v % cvec) * E;
        float ft[E], fx[E];
        unpack_vec<E>(ldg16(tv + v), ft);
        unpack_vec<E>(ldg16(xv + v), fx);
#pragma unroll
        for (int i = 0; i < E; ++i) ft[i] = fmaf(ft[i], p0[cb + i], fmaf(fx[i], p1[cb + i], p2[cb + i]));
        if (res != nullptr) {
            float fr[E];
            unpack_vec<E>(ldg16(rv + v), fr);
#pragma unroll
            for (int i = 0; i < E; ++i) ft[i] += fr[i];
        }
        ov[v] = pack_vec<E>(ft);
    }
}

// ---------------------------------------------------------------------------------------
// ResNet v1 (post-activation) block tail, forward:  y = relu( bn(x) + shortcut )
//   shortcut = none | sc (identity, already activated) | bn_sc(sc) (projection: conv + BN, resnet_model.py:127-168)
// CTA 0 also updates the moving averages of the BatchNorm layers it applies (training).
// ---------------------------------------------------------------------------------------
template <int E>
__global__ void __launch_bounds__(256)
bn_act_apply_kernel(const dtf_bf16* __restrict__ x, DtfBnRef bn, const dtf_bf16* __restrict__ sc, DtfBnRef bn_sc,
                    int sc_mode, dtf_bf16* __restrict__ out, long long total_vec, int C, int update_running) {
    griddep_launch_dependents();
    griddep_wait();
    __shared__ float a0[256], a1[256], b0[256], b1[256];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float mean, invstd;
        bn_consts(bn, c, a0[c], a1[c], mean, invstd);
        b0[c] = 1.f; b1[c] = 0.f;
        if (sc_mode == 2) bn_consts(bn_sc, c, b0[c], b1[c], mean, invstd);
    }
    if (blockIdx.x == 0 && update_running) {
        for (int which = 0; which < (sc_mode == 2 ? 2 : 1); ++which) {
            const DtfBnRef& B = which == 0 ? bn : bn_sc;
            if (B.use_running) continue;
            const float m = 1.f / B.inv_count;
            const float unbias = m > 1.f ? m / (m - 1.f) : 1.f;
            for (int c = threadIdx.x; c < C; c += blockDim.x) {
                const float mean = B.stats[c] * B.inv_count;
                const float var = fmaxf(B.stats[C + c] * B.inv_count - mean * mean, 0.f);
                B.run_mean[c] = B.decay * B.run_mean[c] + (1.f - B.decay) * mean;
                B.run_var[c] = B.decay * B.run_var[c] + (1.f - B.decay) * var * unbias;
            }
        }
    }
    __syncthreads();
    const int cvec = C / E;
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    const uint4* sv = reinterpret_cast<const uint4*>(sc);
    uint4* ov = reinterpret_cast<uint4*>(out);
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total_vec;
         v += (long long)gridDim.x * blockDim.x) {
        const int cb = static_cast<int>(v % cvec) * E;
        float fx[E], fs[E];
        unpack_vec<E>(ldg16(xv + v), fx);
        if (sc_mode != 0) unpack_vec<E>(ldg16(sv + v), fs);
#pragma unroll
        for (int i = 0; i < E; ++i) {
            float y = fmaf(fx[i], a0[cb + i], a1[cb + i]);
            if (sc_mode != 0) y += fmaf(fs[i], b0[cb + i], b1[cb + i]);
            fx[i] = fmaxf(y, 0.f);
        }
        ov[v] = pack_vec<E>(fx);
    }
}

// ---------------------------------------------------------------------------------------
// ResNet v1 block tail, backward:  t = dy * (y > 0), written out, plus the BatchNorm-backward sums of the
// layer(s) that fed the tail:  stats[0..C) += sum t,  stats[C..2C) += sum t * xhat(x)  (= dbeta, dgamma), and the
// same against the projection's BatchNorm input when there is one.  The convolutions upstream then apply the
// BN-backward combination in their operand gather (DTF_PRO_BN_BWD), exactly as in the v2 blocks.
// ---------------------------------------------------------------------------------------
template <int E>
__global__ void __launch_bounds__(256)
relu_bwd_stats_kernel(const dtf_bf16* __restrict__ dy, const dtf_bf16* __restrict__ y, dtf_bf16* __restrict__ t_out,
                      const dtf_bf16* __restrict__ x, DtfBnRef bn, float* __restrict__ stats,
                      const dtf_bf16* __restrict__ x2, DtfBnRef bn2, float* __restrict__ stats2,
                      long long total_vec, int C) {
    griddep_launch_dependents();
    griddep_wait();
    __shared__ float m1[256], i1[256], m2[256], i2[256];
    __shared__ float acc[4][256];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float sc_, sh_;
        bn_consts(bn, c, sc_, sh_, m1[c], i1[c]);
        if (x2 != nullptr) bn_consts(bn2, c, sc_, sh_, m2[c], i2[c]);
        acc[0][c] = acc[1][c] = acc[2][c] = acc[3][c] = 0.f;
    }
    __syncthreads();
    const int cvec = C / E;
    // a thread keeps the same E channels for all its vectors when the stride is a multiple of cvec
    const long long stride = (long long)gridDim.x * blockDim.x;
    const bool fixed = (stride % cvec) == 0;
    float s[4][E];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int i = 0; i < E; ++i) s[k][i] = 0.f;
    int cb_fixed = 0;
    const uint4* dyv = reinterpret_cast<const uint4*>(dy);
    const uint4* yv = reinterpret_cast<const uint4*>(y);
    const uint4* xv = reinterpret_cast<const uint4*>(x);
    const uint4* x2v = reinterpret_cast<const uint4*>(x2);
    uint4* tv = reinterpret_cast<uint4*>(t_out);
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total_vec; v += stride) {
        const int cb = static_cast<int>(v % cvec) * E;
        cb_fixed = cb;
        float g[E], fy[E], fx[E], fx2[E];
        unpack_vec<E>(ldg16(dyv + v), g);
        unpack_vec<E>(ldg16(yv + v), fy);
        unpack_vec<E>(ldg16(xv + v), fx);
        if (x2 != nullptr) unpack_vec<E>(ldg16(x2v + v), fx2);
#pragma unroll
        for (int i = 0; i < E; ++i) {
            const float t = fy[i] > 0.f ? g[i] : 0.f;
            g[i] = t;
            const float h1 = t * (fx[i] - m1[cb + i]) * i1[cb + i];
            const float h2 = x2 != nullptr ? t * (fx2[i] - m2[cb + i]) * i2[cb + i] : 0.f;
            if (fixed) { s[0][i] += t; s[1][i] += h1; s[2][i] += h2; }
            else {
                atomicAdd(&acc[0][cb + i], t); atomicAdd(&acc[1][cb + i], h1);
                if (x2 != nullptr) atomicAdd(&acc[2][cb + i], h2);
            }
        }
        tv[v] = pack_vec<E>(g);
    }
    if (fixed) {
#pragma unroll
        for (int i = 0; i < E; ++i) {
            atomicAdd(&acc[0][cb_fixed + i], s[0][i]);
            atomicAdd(&acc[1][cb_fixed + i], s[1][i]);
            if (x2 != nullptr) atomicAdd(&acc[2][cb_fixed + i], s[2][i]);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        atomicAdd(stats + c, acc[0][c]);
        atomicAdd(stats + C + c, acc[1][c]);
        if (x2 != nullptr) { atomicAdd(stats2 + c, acc[0][c]); atomicAdd(stats2 + C + c, acc[2][c]); }
    }
}

// ---------------------------------------------------------------------------------------
// classifier head: one CTA per image
// ---------------------------------------------------------------------------------------
constexpr int kHeadThreads = 256;
constexpr int kMaxClasses = 16;

__global__ void __launch_bounds__(kHeadThreads)
head_kernel(const __grid_constant__ DtfHead P) {
    griddep_launch_dependents();
    griddep_wait();
    constexpr int kMaxHeadC = 1024;
    __shared__ float s_scale[kMaxHeadC], s_shift[kMaxHeadC], s_mean[kMaxHeadC], s_invstd[kMaxHeadC];
    __shared__ float pooled[kMaxHeadC], gpool[kMaxHeadC], st1[kMaxHeadC], st2[kMaxHeadC];
    __shared__ float logits[kMaxClasses], dlog[kMaxClasses];
    const int n = blockIdx.x;
    const int C = P.C, HW = P.HW, K = P.classes;
    const int tid = threadIdx.x;
    for (int c = tid; c < C; c += kHeadThreads) {
        if (P.has_bn) bn_consts(P.bn, c, s_scale[c], s_shift[c], s_mean[c], s_invstd[c]);
        else { s_scale[c] = 1.f; s_shift[c] = 0.f; s_mean[c] = 0.f; s_invstd[c] = 1.f; }
        pooled[c] = 0.f; st1[c] = 0.f; st2[c] = 0.f;
    }
    if (n == 0 && P.has_bn && P.update_running && !P.bn.use_running) {
        const float m = 1.f / P.bn.inv_count;
        const float unbias = m > 1.f ? m / (m - 1.f) : 1.f;
        for (int c = tid; c < C; c += kHeadThreads) {
            const float mean = P.bn.stats[c] * P.bn.inv_count;
            const float var = fmaxf(P.bn.stats[C + c] * P.bn.inv_count - mean * mean, 0.f);
            P.bn.run_mean[c] = P.bn.decay * P.bn.run_mean[c] + (1.f - P.bn.decay) * mean;
            P.bn.run_var[c] = P.bn.decay * P.bn.run_var[c] + (1.f - P.bn.decay) * var * unbias;
        }
    }
    __syncthreads();
    // element type of x / t_out: bf16, or fp32 for the tf32 engine (P.dtype)
    const bool f32 = P.dtype == DTF_F32;
    const dtf_bf16* xn = P.x + static_cast<size_t>(n) * HW * C * (f32 ? 2 : 1);
    const float* xnf = reinterpret_cast<const float*>(xn);
    // forward: activation + average pool (thread owns channel c = idx % C for all its elements
    // when C divides the thread count, which holds for C in {64, 128, 256})
    if (kHeadThreads % C == 0) {        // a thread keeps one channel for all its elements
        const int c = tid % C;
        float acc = 0.f;
        for (int idx = tid; idx < HW * C; idx += kHeadThreads) {
            const float v = f32 ? xnf[idx] : __bfloat162float(xn[idx]);
            float a = fmaf(v, s_scale[c], s_shift[c]);
            if (P.has_bn) a = fmaxf(a, 0.f);
            acc += a;
        }
        atomicAdd(&pooled[c], acc);
    } else {
        for (int idx = tid; idx < HW * C; idx += kHeadThreads) {
            const int c = idx % C;
            const float v = f32 ? xnf[idx] : __bfloat162float(xn[idx]);
            float a = fmaf(v, s_scale[c], s_shift[c]);
            if (P.has_bn) a = fmaxf(a, 0.f);
            atomicAdd(&pooled[c], a);
        }
    }
    __syncthreads();
    const float inv_hw = 1.f / HW;
    if (tid < K) {
        float z = P.bias[tid];
        for (int c = 0; c < C; ++c) z = fmaf(P.w[tid * C + c], pooled[c] * inv_hw, z);
        logits[tid] = z;
    }
    __syncthreads();
    if (tid == 0) {
        const int label = P.labels[n];
        float mx = logits[0];
        int arg = 0;
        for (int k = 1; k < K; ++k) if (logits[k] > mx) { mx = logits[k]; arg = k; }
        float se = 0.f;
        for (int k = 0; k < K; ++k) se += __expf(logits[k] - mx);
        const float lse = mx + __logf(se);
        const float inv_n = 1.f / P.batch;
        for (int k = 0; k < K; ++k)
            dlog[k] = (__expf(logits[k] - lse) - (k == label ? 1.f : 0.f)) * inv_n * P.loss_scale;
        atomicAdd(P.loss_out, (lse - logits[label]) * inv_n);
        if (arg == label) atomicAdd(P.correct_out, 1);
        // eval-score reduction: the hit also goes straight into this member's slot of the master's score board
        // (a peer pointer when the master's GPU is another one) -- fire-and-forget system-scope reductions
        if (P.board_hits != nullptr) {
            if (arg == label) asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(P.board_hits) : "memory");
            if (n == 0) asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(P.board_hits + 1), "r"(P.batch) : "memory");
        }
    }
    if (P.logits_out != nullptr && tid < K) P.logits_out[n * K + tid] = logits[tid];
    if (!P.training) return;
    __syncthreads();
    // dense gradients
    for (int i = tid; i < K * C; i += kHeadThreads) {
        const int k = i / C, c = i - k * C;
        atomicAdd(P.dw + i, dlog[k] * pooled[c] * inv_hw);
    }
    if (tid < K) atomicAdd(P.db + tid, dlog[tid]);
    for (int c = tid; c < C; c += kHeadThreads) {
        float g = 0.f;
        for (int k = 0; k < K; ++k) g = fmaf(dlog[k], P.w[k * C + c], g);
        gpool[c] = g * inv_hw;
    }
    __syncthreads();
    {
        dtf_bf16* tn = P.t_out + static_cast<size_t>(n) * HW * C * (f32 ? 2 : 1);
        float* tnf = reinterpret_cast<float*>(tn);
        const bool fixed_c = kHeadThreads % C == 0;
        float a1 = 0.f, a2 = 0.f;
        for (int idx = tid; idx < HW * C; idx += kHeadThreads) {
            const int c = fixed_c ? tid % C : idx % C;
            const float v = f32 ? xnf[idx] : __bfloat162float(xn[idx]);
            const bool on = !P.has_bn || fmaf(v, s_scale[c], s_shift[c]) > 0.f;
            const float t = on ? gpool[c] : 0.f;
            if (f32) tnf[idx] = t; else tn[idx] = __float2bfloat16(t);
            if (fixed_c) { a1 += t; a2 += t * (v - s_mean[c]) * s_invstd[c]; }
            else if (P.bwd_stats_out != nullptr) { atomicAdd(&st1[c], t); atomicAdd(&st2[c], t * (v - s_mean[c]) * s_invstd[c]); }
        }
        if (fixed_c) { atomicAdd(&st1[tid % C], a1); atomicAdd(&st2[tid % C], a2); }
    }
    __syncthreads();
    if (P.bwd_stats_out != nullptr)
        for (int c = tid; c < C; c += kHeadThreads) {
            atomicAdd(P.bwd_stats_out + c, st1[c]);
            atomicAdd(P.bwd_stats_out + C + c, st2[c]);
        }
}

// ---------------------------------------------------------------------------------------
// weight packing: fp32 OHWI master -> bf16 [K/8][N][8] for the forward and dgrad GEMMs
// ---------------------------------------------------------------------------------------
DTF_DEVINL void pack_store(const DtfOptSegment& s, dtf_bf16* pack, int local, float w) {
    if (s.pack == 2) {            // dense kernel: plain bf16 copy, row-major as stored
        pack[s.fwd_pack_off + local] = __float2bfloat16(w);
        return;
    }
    // local index inside the OHWI tensor: ((co*taps + tap)*Cin + ci)
    const int ci = local % s.Cin;
    const int rest = local / s.Cin;
    const int tap = rest % s.taps;
    const int co = rest / s.taps;
    if (s.pack == 3) {            // fp32 operands (tf32 MMA): [K/4][N][4] floats, offsets in floats
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
pack_weights_kernel(const float* __restrict__ params, dtf_bf16* __restrict__ pack,
                    const DtfOptSegment* __restrict__ segs, const int2* __restrict__ chunks) {
    const int2 ch = chunks[blockIdx.x];
    const DtfOptSegment s = segs[ch.x];
    if (!s.pack) return;
    for (int i = threadIdx.x; i < DTF_OPT_CHUNK; i += blockDim.x) {
        const int local = ch.y + i;
        if (local < s.count) pack_store(s, pack, local, params[s.offset + local]);
    }
}

// ---------------------------------------------------------------------------------------
// multi-tensor optimizer: TF-1.x Apply{GradientDescent,Momentum,Adagrad,Adadelta,Adam,RMSProp}
// over the flat arena; device-resident hyper-parameters and LR schedule; regulariser gradient
// folded in; optional gradient scale (1/loss_scale, 1/world); emits the bf16 packed weights.
// ---------------------------------------------------------------------------------------
DTF_DEVINL float schedule_lr(const DtfHparams& hp, long long step) {
    // tf.train.piecewise_constant: values[0] for step <= b0, values[i] for b(i-1) < step <= b(i)
    int i = 0;
    while (i < hp.n_boundaries && step > hp.boundaries[i]) ++i;
    return hp.lr_values[i];
}

DTF_DEVINL float apply_rule(const DtfHparams& hp, float lr, float lr_adam, float w, float g, float& s0, float& s1) {
    switch (hp.optimizer) {
        case DTF_OPT_GD:
            return w - lr * g;
        case DTF_OPT_MOMENTUM:
            s0 = hp.momentum * s0 + g;
            return w - lr * s0;
        case DTF_OPT_ADAGRAD:
            s0 += g * g;
            return w - lr * g * rsqrtf(s0);
        case DTF_OPT_ADADELTA: {
            s0 = 0.95f * s0 + 0.05f * g * g;
            const float upd = sqrtf(s1 + 1e-8f) * rsqrtf(s0 + 1e-8f) * g;
            s1 = 0.95f * s1 + 0.05f * upd * upd;
            return w - lr * upd;
        }
        case DTF_OPT_ADAM:
            s0 = 0.9f * s0 + 0.1f * g;
            s1 = 0.999f * s1 + 0.001f * g * g;
            return w - lr_adam * s0 / (sqrtf(s1) + 1e-8f);
        default: {  // DTF_OPT_RMSPROP
            s0 = hp.grad_decay * s0 + (1.f - hp.grad_decay) * g * g;
            s1 = hp.momentum * s1 + lr * g * rsqrtf(s0 + 1e-10f);
            return w - s1;
        }
    }
}

__global__ void __launch_bounds__(256)
optimizer_kernel(float* __restrict__ params, const float* __restrict__ grads,
                 float* __restrict__ slot0, float* __restrict__ slot1, dtf_bf16* __restrict__ pack,
                 const DtfOptSegment* __restrict__ segs, const int2* __restrict__ chunks,
                 const DtfHparams* __restrict__ hp_ptr, const long long* __restrict__ step_ptr,
                 float grad_scale) {
    griddep_launch_dependents();
    griddep_wait();
    __shared__ DtfHparams hp;
    if (threadIdx.x == 0) hp = *hp_ptr;
    __syncthreads();
    const long long step = *step_ptr;             // 1-based count of THIS update (dtf_step_begin bumped it)
    const float lr = schedule_lr(hp, step - 1);   // TF evaluates the schedule at the pre-increment global_step
    float lr_adam = lr;
    if (hp.optimizer == DTF_OPT_ADAM) {
        const float t = static_cast<float>(step);
        lr_adam = lr * sqrtf(1.f - powf(0.999f, t)) / (1.f - powf(0.9f, t));
    }
    const int2 ch = chunks[blockIdx.x];
    const DtfOptSegment s = segs[ch.x];
    const float gs = grad_scale * hp.inv_loss_scale;
    for (int i = threadIdx.x; i < DTF_OPT_CHUNK; i += blockDim.x) {
        const int local = ch.y + i;
        if (local >= s.count) break;
        const int idx = s.offset + local;
        const float w = params[idx];
        float g = grads[idx] * gs;
        if (s.regularized) {
            if (hp.regularizer == DTF_REG_L1 || hp.regularizer == DTF_REG_L1_L2)
                g += hp.weight_decay * (w > 0.f ? 1.f : (w < 0.f ? -1.f : 0.f));
            if (hp.regularizer == DTF_REG_L2 || hp.regularizer == DTF_REG_L1_L2) g += hp.weight_decay * w;
        }
        float s0 = slot0[idx], s1 = slot1[idx];
        const float nw = apply_rule(hp, lr, lr_adam, w, g, s0, s1);
        params[idx] = nw;
        slot0[idx] = s0;
        slot1[idx] = s1;
        if (s.pack) pack_store(s, pack, local, nw);
    }
}

// zero the per-step accumulators (gradient arena, BN statistics, loss/metric cells) and bump
// the device-resident step counter — one launch at the start of every step
__global__ void __launch_bounds__(256)
step_begin_kernel(float4* __restrict__ zero, long long n_vec, long long* __restrict__ step_ptr, int bump) {
    griddep_launch_dependents();
    griddep_wait();
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec;
         i += (long long)gridDim.x * blockDim.x)
        zero[i] = z;
    if (bump && blockIdx.x == 0 && threadIdx.x == 0) *step_ptr += 1;
}

__global__ void fill_kernel(float* __restrict__ p, long long n, float v) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        p[i] = v;
}

// ---------------------------------------------------------------------------------------
// parameter initialisation on the device (reference: the tf.layers initialisers a member is built with --
// glorot_uniform by default, variance_scaling he / glorot-normal as truncated normals, cifar10_main.py:167-199 via
// resnet_model.py:81-92; constants for BatchNorm and biases).  Counter-based Philox4x32-10: element i of segment s
// draws from counter (i / 4, s) under key (seed), so the result depends on (seed, layout) only, not on the launch
// shape.  One thread produces four consecutive elements.
// ---------------------------------------------------------------------------------------
DTF_DEVINL uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}

__global__ void __launch_bounds__(256)
init_params_kernel(float* __restrict__ params, const DtfInitSegment* __restrict__ segs, int n_segs,
                   unsigned long long seed) {
    const int si = blockIdx.y;
    if (si >= n_segs) return;
    const DtfInitSegment s = segs[si];
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    // P(|z| < 2) of a standard normal: the truncated normal is drawn by inverting the CDF on that interval
    const float lo = 0.0227501319f, span = 0.9544997361f;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q * 4 < s.count; q += gridDim.x * blockDim.x) {
        const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(q), 0u, static_cast<uint32_t>(si), 0x0d7fu), key);
        const uint32_t bits[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int i = q * 4 + e;
            if (i >= s.count) break;
            const float u = (static_cast<float>(bits[e] >> 8) + 0.5f) * (1.0f / 16777216.0f);     // (0, 1)
            float v;
            if (s.kind == DTF_INIT_CONST) v = s.scale;
            else if (s.kind == DTF_INIT_UNIFORM) v = (2.f * u - 1.f) * s.scale;
            else v = normcdfinvf(lo + span * u) * s.scale;                                          // DTF_INIT_TRUNC_NORMAL
            // zero the padded input channels of a kernel stored [.., c_pad] with c_real < c_pad real ones
            if (s.c_pad > 0 && (i % s.c_pad) >= s.c_real) v = 0.f;
            params[s.offset + i] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------
// input preprocessing: uint8 NHWC -> bf16 NHWC with channels padded to 8
//   mode 0: plain cast (MNIST: pixels stay 0..255)
//   mode 1: per-image standardisation (CIFAR eval)
//   mode 2: pad 4 + random crop + random flip + standardisation (CIFAR train)
// one CTA per image
// ---------------------------------------------------------------------------------------
DTF_DEVINL uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

template <int E>
__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ images, const int* __restrict__ indices,
                  dtf_bf16* __restrict__ out, const int* __restrict__ crop_flip,
                  int H, int W, int C, int mode, unsigned long long seed, unsigned long long counter,
                  const int* __restrict__ labels_all, int* __restrict__ labels_out,
                  const long long* __restrict__ step) {
    griddep_launch_dependents();
    griddep_wait();
    __shared__ float red[2][8];
    __shared__ float s_mean, s_inv;
    const int n = blockIdx.x;
    const int src_idx = indices ? indices[n] : n;
    if (labels_out != nullptr && threadIdx.x == 0) labels_out[n] = labels_all[src_idx];   // batch gather of the labels
    const uint8_t* img = images + static_cast<size_t>(src_idx) * H * W * C;
    int oy = 4, ox = 4, flip = 0;
    if (mode == 2) {
        if (crop_flip) { oy = crop_flip[3 * n]; ox = crop_flip[3 * n + 1]; flip = crop_flip[3 * n + 2]; }
        else {
            // `step` (device step counter, optional) keeps the draw fresh when the launch is replayed from a CUDA
            // graph, where `counter` is frozen at capture time
            const uint64_t ctr = counter + (step ? static_cast<uint64_t>(*step) * gridDim.x : 0ull);
            const uint64_t r = splitmix64(seed ^ splitmix64(ctr + n));
            oy = static_cast<int>(r % 9); ox = static_cast<int>((r >> 16) % 9); flip = static_cast<int>((r >> 32) & 1);
        }
    }
    const int total = H * W * C;
    float s = 0.f, ss = 0.f;
    if (mode != 0) {
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int c = i % C, p = i / C, x = p % W, y = p / W;
            const int sx = (flip ? (W - 1 - x) : x) + ox - 4, sy = y + oy - 4;
            const float v = (mode == 2 && (sx < 0 || sx >= W || sy < 0 || sy >= H)) ? 0.f
                            : static_cast<float>(img[(sy * W + sx) * C + c]);
            s += v; ss += v * v;
        }
        for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); ss += __shfl_xor_sync(0xffffffffu, ss, o); }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = ss; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float a = 0.f, b = 0.f;
            for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; }
            const float mean = a / total;
            const float var = fmaxf(b / total - mean * mean, 0.f);
            s_mean = mean;
            s_inv = 1.f / fmaxf(sqrtf(var), rsqrtf(static_cast<float>(total)));
        }
        __syncthreads();
    }
    uint4* o = reinterpret_cast<uint4*>(out) + static_cast<size_t>(n) * H * W * (8 / E);   // 8 channels per pixel
    for (int p = threadIdx.x; p < H * W; p += blockDim.x) {
        const int x = p % W, y = p / W;
        const int sx = (mode == 2) ? ((flip ? (W - 1 - x) : x) + ox - 4) : x;
        const int sy = (mode == 2) ? (y + oy - 4) : y;
        const bool inb = sx >= 0 && sx < W && sy >= 0 && sy < H;
        float f[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float v = 0.f;
            if (c < C) {
                v = inb ? static_cast<float>(img[(sy * W + sx) * C + c]) : 0.f;
                if (mode != 0) v = (v - s_mean) * s_inv;
            }
            f[c] = v;
        }
#pragma unroll
        for (int q = 0; q < 8 / E; ++q) o[p * (8 / E) + q] = pack_vec<E>(f + q * E);
    }
}

}  // namespace dtf

// =========================================================================================
// host launchers (C ABI)
// =========================================================================================
#define DTF_CHECK_LAUNCH() (cudaGetLastError() == cudaSuccess ? DTF_OK : DTF_ERR_CUDA)

extern "C" int dtf_bn_bwd_apply_dt(const dtf_bf16* t, const dtf_bf16* x, const dtf_bf16* res, dtf_bf16* out,
                                   const DtfBnRef* bn, const float* bwd_stats, long long pixels, int C, int dtype,
                                   cudaStream_t stream) {
    if (C % 8 != 0 || C > 256) return DTF_ERR_SHAPE;
    if (dtype != DTF_BF16 && dtype != DTF_F32) return DTF_ERR_ARG;
    const long long total_vec = pixels * (C / (dtype == DTF_F32 ? 4 : 8));
    long long blocks = (total_vec + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (dtype == DTF_F32)
        dtf_launch(dtf::bn_bwd_apply_kernel<4>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, t, x, res, out, *bn, bwd_stats, total_vec, C);
    else
        dtf_launch(dtf::bn_bwd_apply_kernel<8>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, t, x, res, out, *bn, bwd_stats, total_vec, C);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_bn_bwd_apply(const dtf_bf16* t, const dtf_bf16* x, const dtf_bf16* res, dtf_bf16* out,
                                const DtfBnRef* bn, const float* bwd_stats, long long pixels, int C,
                                cudaStream_t stream) {
    return dtf_bn_bwd_apply_dt(t, x, res, out, bn, bwd_stats, pixels, C, DTF_BF16, stream);
}

extern "C" int dtf_bn_act_apply_dt(const dtf_bf16* x, const DtfBnRef* bn, const dtf_bf16* sc, const DtfBnRef* bn_sc,
                                   dtf_bf16* out, long long pixels, int C, int update_running, int dtype,
                                   cudaStream_t stream) {
    if (C % 8 != 0 || C > 256 || bn == nullptr) return DTF_ERR_SHAPE;
    if (dtype != DTF_BF16 && dtype != DTF_F32) return DTF_ERR_ARG;
    const long long total_vec = pixels * (C / (dtype == DTF_F32 ? 4 : 8));
    long long blocks = (total_vec + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const int sc_mode = sc == nullptr ? 0 : (bn_sc == nullptr ? 1 : 2);
    DtfBnRef none = {};
    if (dtype == DTF_F32)
        dtf_launch(dtf::bn_act_apply_kernel<4>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, x, *bn, sc,
                   bn_sc ? *bn_sc : none, sc_mode, out, total_vec, C, update_running);
    else
        dtf_launch(dtf::bn_act_apply_kernel<8>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, x, *bn, sc,
                   bn_sc ? *bn_sc : none, sc_mode, out, total_vec, C, update_running);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_bn_act_apply(const dtf_bf16* x, const DtfBnRef* bn, const dtf_bf16* sc, const DtfBnRef* bn_sc,
                                dtf_bf16* out, long long pixels, int C, int update_running, cudaStream_t stream) {
    return dtf_bn_act_apply_dt(x, bn, sc, bn_sc, out, pixels, C, update_running, DTF_BF16, stream);
}

extern "C" int dtf_relu_bwd_stats_dt(const dtf_bf16* dy, const dtf_bf16* y, dtf_bf16* t_out, const dtf_bf16* x,
                                     const DtfBnRef* bn, float* stats, const dtf_bf16* x2, const DtfBnRef* bn2,
                                     float* stats2, long long pixels, int C, int dtype, cudaStream_t stream) {
    if (C % 8 != 0 || C > 256 || bn == nullptr || stats == nullptr) return DTF_ERR_SHAPE;
    if (x2 != nullptr && (bn2 == nullptr || stats2 == nullptr)) return DTF_ERR_ARG;
    if (dtype != DTF_BF16 && dtype != DTF_F32) return DTF_ERR_ARG;
    const int cvec = C / (dtype == DTF_F32 ? 4 : 8);
    const long long total_vec = pixels * cvec;
    long long blocks = (total_vec + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    // keep the grid stride a multiple of the channel-vector count: each thread then owns fixed channels
    while (blocks > 1 && (blocks * 256) % cvec != 0) --blocks;
    DtfBnRef none = {};
    if (dtype == DTF_F32)
        dtf_launch(dtf::relu_bwd_stats_kernel<4>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, dy, y, t_out, x,
                   *bn, stats, x2, bn2 ? *bn2 : none, stats2, total_vec, C);
    else
        dtf_launch(dtf::relu_bwd_stats_kernel<8>, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, dy, y, t_out, x,
                   *bn, stats, x2, bn2 ? *bn2 : none, stats2, total_vec, C);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_relu_bwd_stats(const dtf_bf16* dy, const dtf_bf16* y, dtf_bf16* t_out, const dtf_bf16* x,
                                  const DtfBnRef* bn, float* stats, const dtf_bf16* x2, const DtfBnRef* bn2,
                                  float* stats2, long long pixels, int C, cudaStream_t stream) {
    return dtf_relu_bwd_stats_dt(dy, y, t_out, x, bn, stats, x2, bn2, stats2, pixels, C, DTF_BF16, stream);
}

extern "C" int dtf_head(const DtfHead* p, cudaStream_t stream) {
    if (p->C > 1024 || p->classes > dtf::kMaxClasses) return DTF_ERR_SHAPE;
    dtf_launch(dtf::head_kernel, dim3(p->batch), dim3(dtf::kHeadThreads), 0, stream, *p);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_pack_weights(const float* params, dtf_bf16* pack, const DtfOptSegment* segs,
                                const int* chunks, int n_chunks, cudaStream_t stream) {
    dtf::pack_weights_kernel<<<n_chunks, 256, 0, stream>>>(params, pack, segs, reinterpret_cast<const int2*>(chunks));
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_optimizer_step(float* params, const float* grads, float* slot0, float* slot1, dtf_bf16* pack,
                                  const DtfOptSegment* segs, const int* chunks, int n_chunks,
                                  const DtfHparams* hp, const long long* step, float grad_scale,
                                  cudaStream_t stream) {
    dtf_launch(dtf::optimizer_kernel, dim3(n_chunks), dim3(256), 0, stream, params, grads, slot0, slot1, pack, segs,
               reinterpret_cast<const int2*>(chunks), hp, step, grad_scale);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_step_begin(float* zero_base, long long zero_floats, long long* step, int bump, cudaStream_t stream) {
    if (zero_floats % 4 != 0 || (reinterpret_cast<uintptr_t>(zero_base) & 15) != 0) return DTF_ERR_ARG;
    const long long n_vec = zero_floats / 4;
    long long blocks = (n_vec + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    dtf_launch(dtf::step_begin_kernel, dim3(static_cast<int>(blocks)), dim3(256), 0, stream, reinterpret_cast<float4*>(zero_base), n_vec, step, bump);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_fill(float* p, long long n, float v, cudaStream_t stream) {
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    dtf::fill_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(p, n, v);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_init_params(float* params, const DtfInitSegment* segs_dev, int n_segs, int max_count,
                               unsigned long long seed, cudaStream_t stream) {
    if (n_segs < 1 || max_count < 1 || params == nullptr || segs_dev == nullptr) return DTF_ERR_ARG;
    int gx = (max_count / 4 + 255) / 256;
    if (gx > 64) gx = 64;
    if (gx < 1) gx = 1;
    dtf::init_params_kernel<<<dim3(gx, n_segs), 256, 0, stream>>>(params, segs_dev, n_segs, seed);
    return DTF_CHECK_LAUNCH();
}

extern "C" int dtf_preprocess(const uint8_t* images, const int* indices, dtf_bf16* out, const int* crop_flip,
                              int n, int H, int W, int C, int mode, unsigned long long seed,
                              unsigned long long counter, const long long* step, cudaStream_t stream) {
    if (C > 8 || n < 1) return DTF_ERR_SHAPE;
    const int dtype = mode >> 8;                  // DTF_PP_F32 in the mode word selects fp32 output
    mode &= 0xff;
    if (dtype == DTF_F32)
        dtf_launch(dtf::preprocess_kernel<4>, dim3(n), dim3(256), 0, stream, images, indices, out, crop_flip, H, W, C, mode, seed, counter,
                   static_cast<const int*>(nullptr), static_cast<int*>(nullptr), step);
    else
        dtf_launch(dtf::preprocess_kernel<8>, dim3(n), dim3(256), 0, stream, images, indices, out, crop_flip, H, W, C, mode, seed, counter,
                   static_cast<const int*>(nullptr), static_cast<int*>(nullptr), step);
    return DTF_CHECK_LAUNCH();
}

// Same, reading the batch straight out of a resident dataset: image n comes from images[indices[n]] and its
// label from labels_all[indices[n]] (written to labels_out[n]) -- no separate gather / copy kernels per step.
extern "C" int dtf_preprocess_indexed(const uint8_t* images, const int* indices, const int* labels_all, int* labels_out,
                                      dtf_bf16* out, const int* crop_flip, int n, int H, int W, int C, int mode,
                                      unsigned long long seed, unsigned long long counter, const long long* step,
                                      cudaStream_t stream) {
    if (C > 8 || n < 1 || indices == nullptr || labels_all == nullptr || labels_out == nullptr) return DTF_ERR_ARG;
    const int dtype = mode >> 8;
    mode &= 0xff;
    if (dtype == DTF_F32)
        dtf_launch(dtf::preprocess_kernel<4>, dim3(n), dim3(256), 0, stream, images, indices, out, crop_flip, H, W, C, mode, seed, counter,
                   labels_all, labels_out, step);
    else
        dtf_launch(dtf::preprocess_kernel<8>, dim3(n), dim3(256), 0, stream, images, indices, out, crop_flip, H, W, C, mode, seed, counter,
                   labels_all, labels_out, step);
    return DTF_CHECK_LAUNCH();
}
