// sgd_math.cuh -- the per-element SGD-momentum update shared by sgd.cu and the allreduce epilogues.
#pragma once
#include "common.cuh"

// torch.optim.SGD semantics (reference src/main.py:36: lr 0.01, momentum 0.5):
//   g' = g*grad_scale + weight_decay*p
//   buf = momentum*buf + (1-dampening)*g'          (buf starts at 0, which reproduces buf = g' on step 1)
//   p  -= lr * (nesterov ? g' + momentum*buf : buf)
// __fmul_rn/__fadd_rn stop the compiler from contracting where torch rounds twice.
__device__ __forceinline__ void sgd_update(float& p, float& m, float g, const dlb_sgd_hyper_t& h) {
    g = __fmul_rn(g, h.grad_scale);
    if (h.weight_decay != 0.f) g = fmaf(h.weight_decay, p, g);
    float step = g;
    if (h.momentum != 0.f) {
        m = __fadd_rn(__fmul_rn(m, h.momentum), h.dampening != 0.f ? __fmul_rn(g, 1.f - h.dampening) : g);
        step = h.nesterov ? fmaf(h.momentum, m, g) : m;
    }
    p = fmaf(-h.lr, step, p);
}

// eight consecutive gradient elements as floats (16-byte aligned source)
template <typename GradT>
__device__ __forceinline__ void load_grad8(const GradT* g, float* f);

template <>
__device__ __forceinline__ void load_grad8<float>(const float* g, float* f) {
    const vec16 a = ld_vec(g), b = ld_vec(g + 4);
    VecOf<float>::to_float(a, f);
    VecOf<float>::to_float(b, f + 4);
}
template <>
__device__ __forceinline__ void load_grad8<__nv_bfloat16>(const __nv_bfloat16* g, float* f) {
    VecOf<__nv_bfloat16>::to_float(ld_vec(g), f);
}
template <>
__device__ __forceinline__ void load_grad8<__half>(const __half* g, float* f) {
    VecOf<__half>::to_float(ld_vec(g), f);
}

// Apply the update to the elements of one 16-byte vector of the gradient buffer (VecOf<T>::N of
// them) whose values are already in registers as floats.  `e` is the element index in the flat
// arrays.  Used by the allreduce epilogues, where the reduced gradient never leaves registers
// (one-shot) or was just written by the peers into local memory (two-shot).
template <int N>
__device__ __forceinline__ void sgd_apply_vec(const dlb_fused_sgd_t& s, size_t e, const float* g) {
    float pf[N], mf[N];
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
        VecOf<float>::to_float(ld_vec(s.param + e + 4 * i), pf + 4 * i);
        VecOf<float>::to_float(ld_vec(s.momentum_buf + e + 4 * i), mf + 4 * i);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) sgd_update(pf[i], mf[i], g[i], s.hyper);
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
        st_vec(s.param + e + 4 * i, VecOf<float>::from_float(pf + 4 * i));
        st_vec(s.momentum_buf + e + 4 * i, VecOf<float>::from_float(mf + 4 * i));
    }
    if (s.shadow_bf16) {
        __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(s.shadow_bf16) + e;
        if (N == 8) {
            st_vec(sh, VecOf<__nv_bfloat16>::from_float(pf));
        } else {
            // four bf16 = 8 bytes
            __nv_bfloat162 a = __floats2bfloat162_rn(pf[0], pf[1]), b = __floats2bfloat162_rn(pf[2], pf[3]);
            uint2 w;
            w.x = *reinterpret_cast<uint32_t*>(&a);
            w.y = *reinterpret_cast<uint32_t*>(&b);
            *reinterpret_cast<uint2*>(sh) = w;
        }
    }
}
