// sgd.cu -- K-sgd: fused SGD-momentum update that consumes the reduced gradient buffer in place.
//
// Replaces torch.optim.SGD(lr=0.01, momentum=0.5) of reference src/main.py:36,91 (a chain of
// foreach mul/add kernels over all parameters) by one pass: read p, m, g -- write p, m and,
// for bf16 models, the bf16 copy of p.  The gradient may be the fp32 or bf16 fusion buffer
// exactly as the allreduce left it, so no unpack is needed in between.  Rounding follows
// torch's sequence (buf*momentum rounded, then + g; p = fma(-lr, buf, p)) so results agree with
// torch.optim.SGD to the last bit for dampening = 0.
//
// The same per-vector update (`sgd_apply_vec`) is the fused epilogue of the allreduce kernels.
#include "sgd_math.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 8192;       // elements per CTA
constexpr int kMaxPerLaunch = 64;  // tensors per launch of the multi-tensor form

struct SgdTable {
    float* param[kMaxPerLaunch];
    float* mom[kMaxPerLaunch];
    const void* grad[kMaxPerLaunch];
    void* shadow[kMaxPerLaunch];
    uint32_t numel[kMaxPerLaunch];
    uint32_t chunk_prefix[kMaxPerLaunch + 1];
    uint8_t gdtype[kMaxPerLaunch];
    int ntensors;
};

template <typename GradT>
__device__ __forceinline__ void sgd_chunk(float* __restrict__ p, float* __restrict__ m, const GradT* __restrict__ g,
                                          __nv_bfloat16* __restrict__ shadow, int cnt, const dlb_sgd_hyper_t& h) {
    constexpr int G = 8;  // elements per thread-iteration: two fp32 vectors, one 16-bit vector
    const int tid = threadIdx.x;
    const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) |
                           reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(shadow)) & 15) == 0;
    int done = 0;
    if (aligned) {
        const int groups = cnt / G;
        for (int gi = tid; gi < groups; gi += kThreads) {
            const size_t e = (size_t)gi * G;
            float pf[G], mf[G], gf[G];
            // all loads first: 6 (fp32 grad) or 5 (16-bit grad) independent 16-byte requests
            const vec16 p0 = ld_vec(p + e), p1 = ld_vec(p + e + 4);
            const vec16 m0 = ld_vec(m + e), m1 = ld_vec(m + e + 4);
            load_grad8<GradT>(g + e, gf);
            VecOf<float>::to_float(p0, pf);
            VecOf<float>::to_float(p1, pf + 4);
            VecOf<float>::to_float(m0, mf);
            VecOf<float>::to_float(m1, mf + 4);
#pragma unroll
            for (int i = 0; i < G; ++i) sgd_update(pf[i], mf[i], gf[i], h);
            st_vec(p + e, VecOf<float>::from_float(pf));
            st_vec(p + e + 4, VecOf<float>::from_float(pf + 4));
            st_vec(m + e, VecOf<float>::from_float(mf));
            st_vec(m + e + 4, VecOf<float>::from_float(mf + 4));
            if (shadow) st_vec(shadow + e, VecOf<__nv_bfloat16>::from_float(pf));
        }
        done = groups * G;
    }
    for (int i = done + tid; i < cnt; i += kThreads) {
        float pf = p[i], mf = m[i];
        sgd_update(pf, mf, VecOf<GradT>::load1(g + i), h);
        p[i] = pf;
        m[i] = mf;
        if (shadow) shadow[i] = __float2bfloat16_rn(pf);
    }
}

__global__ void __launch_bounds__(kThreads) sgd_multi_kernel(const __grid_constant__ SgdTable tab,
                                                             const __grid_constant__ dlb_sgd_hyper_t h) {
    const uint32_t b = blockIdx.x;
    int lo = 0, hi = tab.ntensors;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tab.chunk_prefix[mid] <= b) lo = mid; else hi = mid;
    }
    const int t = lo;
    const size_t start = (size_t)(b - tab.chunk_prefix[t]) * kChunk;
    const int cnt = (int)min((size_t)kChunk, (size_t)tab.numel[t] - start);
    float* p = tab.param[t] + start;
    float* m = tab.mom[t] + start;
    __nv_bfloat16* sh = tab.shadow[t] ? reinterpret_cast<__nv_bfloat16*>(tab.shadow[t]) + start : nullptr;
    switch (tab.gdtype[t]) {
        case DLB_F32: sgd_chunk<float>(p, m, reinterpret_cast<const float*>(tab.grad[t]) + start, sh, cnt, h); break;
        case DLB_BF16:
            sgd_chunk<__nv_bfloat16>(p, m, reinterpret_cast<const __nv_bfloat16*>(tab.grad[t]) + start, sh, cnt, h);
            break;
        default: sgd_chunk<__half>(p, m, reinterpret_cast<const __half*>(tab.grad[t]) + start, sh, cnt, h); break;
    }
}

// Flat form: grid-stride over chunks so a 100 MiB buffer does not need 100k CTAs' worth of
// launch overhead; 8 CTAs per SM keep ~6 x 16 B x 256 x 8 = 196 KB in flight per SM.
template <typename GradT>
__global__ void __launch_bounds__(kThreads) sgd_flat_kernel(float* __restrict__ p, float* __restrict__ m,
                                                            const GradT* __restrict__ g,
                                                            __nv_bfloat16* __restrict__ shadow, int64_t numel,
                                                            const __grid_constant__ dlb_sgd_hyper_t h) {
    const int64_t nchunks = (numel + kChunk - 1) / kChunk;
    for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const int64_t start = c * kChunk;
        const int cnt = (int)min((int64_t)kChunk, numel - start);
        sgd_chunk<GradT>(p + start, m + start, g + start, shadow ? shadow + start : nullptr, cnt, h);
    }
}

int sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            sms = 148;
    }
    return sms;
}

}  // namespace

extern "C" int dlb_sgd_flat(float* param, float* momentum_buf, const void* grad, int grad_dtype, void* shadow_bf16,
                            int64_t numel, const dlb_sgd_hyper_t* hyper, void* stream_) {
    DLB_REQUIRE(param && momentum_buf && grad && hyper, DLB_E_ARG, "dlb_sgd_flat: null argument");
    DLB_REQUIRE(grad_dtype >= DLB_F32 && grad_dtype <= DLB_F16, DLB_E_ARG, "dlb_sgd_flat: bad grad dtype");
    if (numel <= 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t nchunks = (numel + kChunk - 1) / kChunk;
    const int64_t cap = (int64_t)sm_count() * 8;
    const int grid = (int)(nchunks < cap ? nchunks : cap);
    __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(shadow_bf16);
    switch (grad_dtype) {
        case DLB_F32:
            sgd_flat_kernel<float><<<grid, kThreads, 0, stream>>>(param, momentum_buf, (const float*)grad, sh, numel, *hyper);
            break;
        case DLB_BF16:
            sgd_flat_kernel<__nv_bfloat16>
                <<<grid, kThreads, 0, stream>>>(param, momentum_buf, (const __nv_bfloat16*)grad, sh, numel, *hyper);
            break;
        default:
            sgd_flat_kernel<__half><<<grid, kThreads, 0, stream>>>(param, momentum_buf, (const __half*)grad, sh, numel, *hyper);
            break;
    }
    DLB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int dlb_sgd_multi(void* const* params, void* const* momentum_bufs, void* const* grads,
                             const int* grad_dtypes, const int64_t* numels, int ntensors,
                             const dlb_sgd_hyper_t* hyper, void* stream_) {
    DLB_REQUIRE(ntensors >= 0 && hyper, DLB_E_ARG, "dlb_sgd_multi: bad argument");
    cudaStream_t stream = (cudaStream_t)stream_;
    for (int base = 0; base < ntensors; base += kMaxPerLaunch) {
        const int n = ntensors - base < kMaxPerLaunch ? ntensors - base : kMaxPerLaunch;
        SgdTable tab;
        uint32_t chunks = 0;
        int k = 0;
        for (int i = 0; i < n; ++i) {
            const int j = base + i;
            if (grads[j] == nullptr || numels[j] == 0) continue;  // parameter without gradient: untouched, as in torch
            DLB_REQUIRE(numels[j] > 0 && numels[j] < (int64_t)UINT32_MAX, DLB_E_LIMIT, "tensor %d: numel beyond 2^32-1", j);
            DLB_REQUIRE(grad_dtypes[j] >= DLB_F32 && grad_dtypes[j] <= DLB_F16, DLB_E_ARG, "tensor %d: bad dtype", j);
            tab.param[k] = (float*)params[j];
            tab.mom[k] = (float*)momentum_bufs[j];
            tab.grad[k] = grads[j];
            tab.shadow[k] = nullptr;
            tab.numel[k] = (uint32_t)numels[j];
            tab.gdtype[k] = (uint8_t)grad_dtypes[j];
            tab.chunk_prefix[k] = chunks;
            chunks += (uint32_t)((numels[j] + kChunk - 1) / kChunk);
            ++k;
        }
        if (k == 0) continue;
        tab.chunk_prefix[k] = chunks;
        tab.ntensors = k;
        sgd_multi_kernel<<<chunks, kThreads, 0, stream>>>(tab, *hyper);
        DLB_CUDA(cudaGetLastError());
    }
    return 0;
}
