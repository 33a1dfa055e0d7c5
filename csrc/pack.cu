// pack.cu -- K-pack / K-unpack: multi-tensor gather(+scale+cast) into a flat fusion buffer
// and the reverse scatter.
//
// Replaces reference Group.fuse / Group.unfuse (src/ourdist.py:28-43), which issue one ATen copy
// per tensor (187 copies per step for GoogLeNet, each a blocking D2H when the buffer is on the
// host) followed by a separate full-buffer divide (src/allreduce.py:98).  Here one launch
// handles up to DLB_MAX_TENSORS tensors: the tensor table travels in the kernel parameters (no
// device-side table, nothing to allocate or copy), each CTA finds its (tensor, chunk) by binary
// search in the chunk prefix sums, and scale + dtype conversion happen in registers between a
// 16-byte load and a 16-byte store.  Bandwidth bound: bytes = numel * (src + dst element size).
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 8192;  // elements per CTA
constexpr int kUnroll = 8;    // 16-byte groups in flight per thread: a CTA issues all its loads up front

struct TensorTable {
    void* tensor[DLB_MAX_TENSORS];
    uint32_t offset[DLB_MAX_TENSORS];  // element offset in the flat buffer
    uint32_t numel[DLB_MAX_TENSORS];
    uint32_t chunk_prefix[DLB_MAX_TENSORS + 1];
    uint8_t dtype[DLB_MAX_TENSORS];
    int ntensors;
};

// Copy `cnt` elements src -> dst with scale and conversion.  `src == nullptr` writes zeros.
template <typename SrcT, typename DstT>
__device__ __forceinline__ void copy_chunk(const SrcT* __restrict__ src, DstT* __restrict__ dst, int cnt, float scale) {
    constexpr int NS = VecOf<SrcT>::N, ND = VecOf<DstT>::N;
    constexpr int G = NS > ND ? NS : ND;
    const int tid = threadIdx.x;
    if (src == nullptr) {
        for (int i = tid; i < cnt; i += kThreads) VecOf<DstT>::store1(dst + i, 0.f);
        return;
    }
    const bool aligned = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
    int done = 0;
    if (aligned) {
        const int groups = cnt / G;
        // kUnroll independent 16-byte loads are issued before the first store
        for (int g0 = 0; g0 < groups; g0 += kThreads * kUnroll) {
            float f[kUnroll][G];
            bool live[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const int g = g0 + u * kThreads + tid;
                live[u] = g < groups;
                if (live[u]) {
#pragma unroll
                    for (int i = 0; i < G / NS; ++i)
                        VecOf<SrcT>::to_float(ld_vec_stream(src + (size_t)g * G + i * NS), f[u] + i * NS);
                }
            }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                if (live[u]) {
                    const int g = g0 + u * kThreads + tid;
#pragma unroll
                    for (int i = 0; i < G; ++i) f[u][i] *= scale;
#pragma unroll
                    for (int i = 0; i < G / ND; ++i)
                        st_vec(dst + (size_t)g * G + i * ND, VecOf<DstT>::from_float(f[u] + i * ND));
                }
            }
        }
        done = groups * G;
    }
    for (int i = done + tid; i < cnt; i += kThreads) VecOf<DstT>::store1(dst + i, VecOf<SrcT>::load1(src + i) * scale);
}

template <typename FlatT, typename TensorT, bool PACK>
__device__ __forceinline__ void move_chunk(void* tensor, FlatT* flat, int start, int cnt, float scale) {
    TensorT* t = tensor ? reinterpret_cast<TensorT*>(tensor) + start : nullptr;
    if (PACK) {
        copy_chunk<TensorT, FlatT>(t, flat + start, cnt, scale);
    } else if (t != nullptr) {
        copy_chunk<FlatT, TensorT>(flat + start, t, cnt, scale);
    }
}

template <typename FlatT, bool PACK>
__global__ void __launch_bounds__(kThreads) multi_tensor_move_kernel(const __grid_constant__ TensorTable tab,
                                                                     FlatT* __restrict__ flat, float scale) {
    const uint32_t b = blockIdx.x;
    int lo = 0, hi = tab.ntensors;  // invariant: chunk_prefix[lo] <= b < chunk_prefix[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tab.chunk_prefix[mid] <= b) lo = mid; else hi = mid;
    }
    const int t = lo;
    const uint32_t start = (b - tab.chunk_prefix[t]) * (uint32_t)kChunk;
    const uint32_t n = tab.numel[t];
    const int cnt = (int)min((uint32_t)kChunk, n - start);
    FlatT* dst = flat + tab.offset[t];
    void* tensor = tab.tensor[t];
    switch (tab.dtype[t]) {
        case DLB_F32: move_chunk<FlatT, float, PACK>(tensor, dst, start, cnt, scale); break;
        case DLB_BF16: move_chunk<FlatT, __nv_bfloat16, PACK>(tensor, dst, start, cnt, scale); break;
        default: move_chunk<FlatT, __half, PACK>(tensor, dst, start, cnt, scale); break;
    }
}

template <bool PACK>
int launch_move(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes, const int64_t* offsets,
                const int64_t* numels, int ntensors, float scale, cudaStream_t stream) {
    DLB_REQUIRE(ntensors >= 0, DLB_E_ARG, "dlb_%s: negative tensor count", PACK ? "pack" : "unpack");
    DLB_REQUIRE(flat_dtype == DLB_F32 || flat_dtype == DLB_BF16, DLB_E_ARG, "flat buffer dtype must be fp32 or bf16");
    // lists longer than one table are split into several launches
    for (int base = 0; base < ntensors; base += DLB_MAX_TENSORS) {
        const int n = ntensors - base < DLB_MAX_TENSORS ? ntensors - base : DLB_MAX_TENSORS;
        TensorTable tab;
        uint32_t chunks = 0;
        for (int i = 0; i < n; ++i) {
            const int j = base + i;
            DLB_REQUIRE(numels[j] >= 0 && numels[j] < (int64_t)UINT32_MAX && offsets[j] >= 0 &&
                            offsets[j] < (int64_t)UINT32_MAX,
                        DLB_E_LIMIT, "tensor %d: numel/offset beyond 2^32-1 elements", j);
            DLB_REQUIRE(tensor_dtypes[j] >= DLB_F32 && tensor_dtypes[j] <= DLB_F16, DLB_E_ARG, "tensor %d: bad dtype", j);
            tab.tensor[i] = tensors[j];
            tab.offset[i] = (uint32_t)offsets[j];
            tab.numel[i] = (uint32_t)numels[j];
            tab.dtype[i] = (uint8_t)tensor_dtypes[j];
            tab.chunk_prefix[i] = chunks;
            chunks += (uint32_t)((numels[j] + kChunk - 1) / kChunk);
        }
        tab.chunk_prefix[n] = chunks;
        tab.ntensors = n;
        if (chunks == 0) continue;
        if (flat_dtype == DLB_F32)
            multi_tensor_move_kernel<float, PACK><<<chunks, kThreads, 0, stream>>>(tab, (float*)flat, scale);
        else
            multi_tensor_move_kernel<__nv_bfloat16, PACK><<<chunks, kThreads, 0, stream>>>(tab, (__nv_bfloat16*)flat, scale);
        DLB_CUDA(cudaGetLastError());
    }
    return 0;
}

}  // namespace

extern "C" int dlb_pack(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                        const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream) {
    return launch_move<true>(flat, flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors, scale,
                             (cudaStream_t)stream);
}

extern "C" int dlb_unpack(const void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                          const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream) {
    return launch_move<false>(const_cast<void*>(flat), flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors,
                              scale, (cudaStream_t)stream);
}
