// common.cuh -- shared device helpers for libdlb200 (sm_100a).
//
// 16-byte vector access, bf16/f16 <-> fp32 conversion of a vector, system-scope flag
// accesses for the cross-GPU barrier, and the host-side error plumbing.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/dlb200.h"

// ---- host-side error plumbing ---------------------------------------------------------------
void dlb_set_error(const char* fmt, ...);

#define DLB_CUDA(call)                                                                     \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            dlb_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return (int)e__;                                                               \
        }                                                                                  \
    } while (0)

#define DLB_REQUIRE(cond, code, ...)  \
    do {                              \
        if (!(cond)) {                \
            dlb_set_error(__VA_ARGS__); \
            return (code);            \
        }                             \
    } while (0)

static inline int dlb_dtype_size(int dtype) { return dtype == DLB_F32 ? 4 : 2; }

// ---- 16-byte vectors ------------------------------------------------------------------------
struct __align__(16) vec16 {
    uint32_t x, y, z, w;
};

// plain (weak) global accesses; used for local memory and for peer memory after an acquire
__device__ __forceinline__ vec16 ld_vec(const void* p) {
    vec16 v;
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
// streaming load that does not allocate in L1 (data touched once)
__device__ __forceinline__ vec16 ld_vec_stream(const void* p) {
    vec16 v;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
// read-only path for data that no one writes during the kernel
__device__ __forceinline__ vec16 ld_vec_nc(const void* p) {
    vec16 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st_vec(void* p, const vec16& v) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_vec_stream(void* p, const vec16& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

// ---- element type traits: a vec16 holds VEC<T>::N elements ---------------------------------------
template <typename T>
struct VecOf;

template <>
struct VecOf<float> {
    static constexpr int N = 4;
    __device__ __forceinline__ static void to_float(const vec16& v, float* f) {
        f[0] = __uint_as_float(v.x);
        f[1] = __uint_as_float(v.y);
        f[2] = __uint_as_float(v.z);
        f[3] = __uint_as_float(v.w);
    }
    __device__ __forceinline__ static vec16 from_float(const float* f) {
        vec16 v;
        v.x = __float_as_uint(f[0]);
        v.y = __float_as_uint(f[1]);
        v.z = __float_as_uint(f[2]);
        v.w = __float_as_uint(f[3]);
        return v;
    }
    __device__ __forceinline__ static float load1(const float* p) { return *p; }
    __device__ __forceinline__ static void store1(float* p, float x) { *p = x; }
};

template <>
struct VecOf<__nv_bfloat16> {
    static constexpr int N = 8;
    __device__ __forceinline__ static void to_float(const vec16& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // a bf16 is the upper half of the fp32 with the same value
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ __forceinline__ static vec16 from_float(const float* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        vec16 v;
        v.x = w[0];
        v.y = w[1];
        v.z = w[2];
        v.w = w[3];
        return v;
    }
    __device__ __forceinline__ static float load1(const __nv_bfloat16* p) { return __bfloat162float(*p); }
    __device__ __forceinline__ static void store1(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }
};

template <>
struct VecOf<__half> {
    static constexpr int N = 8;
    __device__ __forceinline__ static void to_float(const vec16& v, float* f) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 p = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            f[2 * i] = p.x;
            f[2 * i + 1] = p.y;
        }
    }
    __device__ __forceinline__ static vec16 from_float(const float* f) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        vec16 v;
        v.x = w[0];
        v.y = w[1];
        v.z = w[2];
        v.w = w[3];
        return v;
    }
    __device__ __forceinline__ static float load1(const __half* p) { return __half2float(*p); }
    __device__ __forceinline__ static void store1(__half* p, float x) { *p = __float2half_rn(x); }
};

// ---- system-scope flag accesses (cross-GPU barrier) --------------------------------------------
__device__ __forceinline__ void st_flag_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_flag_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
