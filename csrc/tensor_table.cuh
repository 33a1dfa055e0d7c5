// tensor_table.cuh -- the list of tensors behind one fusion buffer, as the kernels see it.
//
// A fusion bucket is a flat buffer plus, for every tensor in it, a pointer, an element offset
// and an element count (reference Group, src/ourdist.py:13-51).  The table travels in the
// kernel parameters -- gradient tensors live at new addresses every step, so there is nothing
// worth keeping in device memory -- and each CTA stages it in shared memory once.  The kernels
// are driven by *flat position*: for 16-byte vector v of the flat buffer, `table_find` gives the
// tensor that owns it (binary search in shared memory, with the previous answer as a hint) and
// `gather_vec` / `scatter_vec` move that vector between the tensor's own storage and registers,
// converting between fp32 / bf16 / fp16 on the way.  Work is therefore split in equal pieces of
// the flat buffer whatever the tensor sizes are (GoogLeNet's bucket holds 4-byte biases next to
// 4 MB weights), and the same code is the stand-alone pack/unpack kernel (pack.cu) and the
// prologue of the allreduce kernels (allreduce.cu), where a CTA packs exactly the tiles its
// peers' same-numbered CTAs will pull.
//
// Layout contract for this path ("vector-aligned"): offsets ascend, every tensor starts on a
// 16-byte boundary of the flat buffer, and tensor regions do not overlap.  The host checks it
// and falls back to the element-wise kernel otherwise.
#pragma once
#include "common.cuh"

struct TensorTable {
    void* tensor[DLB_MAX_TENSORS];
    uint32_t offset[DLB_MAX_TENSORS];  // element offset in the flat buffer
    uint32_t numel[DLB_MAX_TENSORS];
    uint8_t dtype[DLB_MAX_TENSORS];
    int ntensors;
    int emu_stacked;  // single-GPU emulation: tensor i of emulated rank r is at tensor[i] + r * numel[i] * element size
};

struct TableSmem {
    void* tensor[DLB_MAX_TENSORS];
    uint32_t offset[DLB_MAX_TENSORS];
    uint32_t numel[DLB_MAX_TENSORS];
    uint8_t dtype[DLB_MAX_TENSORS];
};

__device__ __forceinline__ int dlb_esize(int dtype) { return dtype == DLB_F32 ? 4 : 2; }

// every thread of the CTA calls this; ends with a CTA barrier
__device__ __forceinline__ void table_stage(TableSmem& s, const TensorTable& tab, int emu_rank) {
    for (int i = threadIdx.x; i < tab.ntensors; i += blockDim.x) {
        char* p = reinterpret_cast<char*>(tab.tensor[i]);
        const uint32_t n = tab.numel[i];
        const uint8_t dt = tab.dtype[i];
        if (p != nullptr && tab.emu_stacked) p += (size_t)emu_rank * n * dlb_esize(dt);
        s.tensor[i] = p;
        s.offset[i] = tab.offset[i];
        s.numel[i] = n;
        s.dtype[i] = dt;
    }
    __syncthreads();
}

// last tensor whose offset is <= e (0 when e lies before the first tensor)
__device__ __forceinline__ int table_find(const TableSmem& s, int n, uint32_t e, int hint) {
    if (hint >= 0 && s.offset[hint] <= e && (hint + 1 == n || e < s.offset[hint + 1])) return hint;
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s.offset[mid] <= e) lo = mid; else hi = mid;
    }
    return lo;
}

// N consecutive elements wanted from `src`, `cnt` (<= N) of them exist; missing ones read as zero.
// N * sizeof(SrcT) is 8, 16 or 32 bytes: one 8-byte load, or one or two 16-byte loads when aligned.
template <typename SrcT, int N>
__device__ __forceinline__ void load_elems(const SrcT* __restrict__ src, int cnt, float* f) {
    constexpr int NS = VecOf<SrcT>::N;
    constexpr int BYTES = N * (int)sizeof(SrcT);
    const uintptr_t a = reinterpret_cast<uintptr_t>(src);
    if (cnt == N && (a & (BYTES >= 16 ? 15 : 7)) == 0) {
        if constexpr (BYTES >= 16) {
#pragma unroll
            for (int i = 0; i < BYTES / 16; ++i) VecOf<SrcT>::to_float(ld_vec_stream(src + i * NS), f + i * NS);
        } else {
            uint2 w;
            asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w.x), "=r"(w.y) : "l"(src));
            vec16 v;
            v.x = w.x;
            v.y = w.y;
            v.z = v.w = 0;
            float g[NS];
            VecOf<SrcT>::to_float(v, g);
#pragma unroll
            for (int i = 0; i < N; ++i) f[i] = g[i];
        }
    } else {
#pragma unroll
        for (int i = 0; i < N; ++i) f[i] = i < cnt ? VecOf<SrcT>::load1(src + i) : 0.f;
    }
}

// the first `cnt` (<= N) of N floats go to `dst`
template <typename DstT, int N>
__device__ __forceinline__ void store_elems(DstT* __restrict__ dst, int cnt, const float* f) {
    constexpr int ND = VecOf<DstT>::N;
    constexpr int BYTES = N * (int)sizeof(DstT);
    const uintptr_t a = reinterpret_cast<uintptr_t>(dst);
    if (cnt == N && (a & (BYTES >= 16 ? 15 : 7)) == 0) {
        if constexpr (BYTES >= 16) {
#pragma unroll
            for (int i = 0; i < BYTES / 16; ++i) st_vec(dst + i * ND, VecOf<DstT>::from_float(f + i * ND));
        } else {
            float g[ND];
#pragma unroll
            for (int i = 0; i < ND; ++i) g[i] = i < N ? f[i] : 0.f;
            const vec16 v = VecOf<DstT>::from_float(g);
            *reinterpret_cast<uint2*>(dst) = make_uint2(v.x, v.y);
        }
    } else {
#pragma unroll
        for (int i = 0; i < N; ++i)
            if (i < cnt) VecOf<DstT>::store1(dst + i, f[i]);
    }
}

// The tensor a thread is currently working in, kept in registers: consecutive vectors of a thread
// almost always belong to the same tensor, so the shared-memory table is only consulted when the
// thread crosses into another one.
struct Owner {
    uint32_t lo, hi;  // flat elements [lo, hi) are this tensor's
    char* base;       // its storage (nullptr: none)
    int dtype, t;
    __device__ __forceinline__ Owner() : lo(1), hi(0), base(nullptr), dtype(0), t(-1) {}
};

// true when flat element `e` belongs to a tensor (then `own` describes it); false for a hole
__device__ __forceinline__ bool owner_of(const TableSmem& s, int n, uint32_t e, Owner& own) {
    if (e >= own.lo && e < own.hi) return true;
    const int t = table_find(s, n, e, own.t);
    own.t = t;
    own.lo = s.offset[t];
    own.hi = own.lo + s.numel[t];
    own.base = reinterpret_cast<char*>(s.tensor[t]);
    own.dtype = s.dtype[t];
    return e >= own.lo && e < own.hi;
}

// The N = VecOf<FlatT>::N elements of flat vector `v`, fetched from the tensor that owns them, in
// two steps so that a thread can have several vectors in flight: `gather_issue` finds the owner and
// issues the 16-byte load(s) without touching the result, `gather_finish` converts to floats.
// A vector that belongs to no tensor (a hole in the layout) finishes as `false`: nothing to write.
// Elements past the tensor's end (the pad up to the next 16-byte boundary) and tensors without
// storage (a parameter without gradient) read as zero.
struct Gathered {
    vec16 w0, w1;  // raw source bytes (w1: second half of 32 source bytes, fp32 source for a 16-bit flat)
    int kind;      // -1 hole | 0 zeros | 1 fp32 raw | 2 bf16 raw | 3 fp16 raw | 4 floats (slow path, already converted)
};

template <typename FlatT>
__device__ __forceinline__ void gather_issue(const TableSmem& s, int n, uint32_t v, Owner& own, Gathered& g) {
    constexpr int N = VecOf<FlatT>::N;
    const uint32_t e = v * N;
    if (!owner_of(s, n, e, own)) {
        g.kind = -1;
        return;
    }
    if (own.base == nullptr) {
        g.kind = 0;
        return;
    }
    const uint32_t k = e - own.lo;
    const int cnt = (int)min((uint32_t)N, own.hi - e);
    const int dt = own.dtype;
    const int es = dlb_esize(dt);
    const char* src = own.base + (size_t)k * es;
    const int bytes = N * es;  // 8, 16 or 32
    if (cnt == N && (reinterpret_cast<uintptr_t>(src) & (bytes >= 16 ? 15 : 7)) == 0) {
        g.kind = dt == DLB_F32 ? 1 : (dt == DLB_BF16 ? 2 : 3);
        if (bytes == 8) {
            asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(g.w0.x), "=r"(g.w0.y) : "l"(src));
        } else {
            g.w0 = ld_vec_stream(src);
            if (bytes == 32) g.w1 = ld_vec_stream(src + 16);
        }
        return;
    }
    float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    switch (dt) {
        case DLB_F32: load_elems<float, N>(reinterpret_cast<const float*>(src), cnt, f); break;
        case DLB_BF16: load_elems<__nv_bfloat16, N>(reinterpret_cast<const __nv_bfloat16*>(src), cnt, f); break;
        default: load_elems<__half, N>(reinterpret_cast<const __half*>(src), cnt, f); break;
    }
    g.w0 = VecOf<float>::from_float(f);
    g.w1 = VecOf<float>::from_float(f + 4);
    g.kind = 4;
}

template <typename FlatT>
__device__ __forceinline__ bool gather_finish(const Gathered& g, float* f) {
    constexpr int N = VecOf<FlatT>::N;
    if (g.kind < 0) return false;
    if (g.kind == 1 || g.kind == 4) {
        VecOf<float>::to_float(g.w0, f);
        if (N == 8) VecOf<float>::to_float(g.w1, f + (N == 8 ? 4 : 0));
    } else if (g.kind == 2) {
        float t[8];
        VecOf<__nv_bfloat16>::to_float(g.w0, t);
#pragma unroll
        for (int i = 0; i < N; ++i) f[i] = t[i];
    } else if (g.kind == 3) {
        float t[8];
        VecOf<__half>::to_float(g.w0, t);
#pragma unroll
        for (int i = 0; i < N; ++i) f[i] = t[i];
    } else {
#pragma unroll
        for (int i = 0; i < N; ++i) f[i] = 0.f;
    }
    return true;
}

// The reverse: the N elements of flat vector `v` (already floats) go to the tensor that owns them.
template <typename FlatT>
__device__ __forceinline__ void scatter_vec(const TableSmem& s, int n, uint32_t v, Owner& own, const float* f) {
    constexpr int N = VecOf<FlatT>::N;
    const uint32_t e = v * N;
    if (!owner_of(s, n, e, own) || own.base == nullptr) return;
    const uint32_t k = e - own.lo;
    const int cnt = (int)min((uint32_t)N, own.hi - e);
    switch (own.dtype) {
        case DLB_F32: store_elems<float, N>(reinterpret_cast<float*>(own.base) + k, cnt, f); break;
        case DLB_BF16: store_elems<__nv_bfloat16, N>(reinterpret_cast<__nv_bfloat16*>(own.base) + k, cnt, f); break;
        default: store_elems<__half, N>(reinterpret_cast<__half*>(own.base) + k, cnt, f); break;
    }
}

// ---- host side: fill a table from the C-ABI arrays -------------------------------------------------
// Returns 0, or DLB_E_* with the error text set.  `vector_aligned` reports whether the layout
// satisfies the contract above for a flat buffer of `flat_dtype`.
int dlb_fill_table(TensorTable& tab, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                   const int64_t* offsets, const int64_t* numels, int base, int n, bool* vector_aligned,
                   uint32_t* first_elem, uint32_t* end_elem);
