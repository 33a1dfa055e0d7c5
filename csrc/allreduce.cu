// allreduce.cu -- K-ar: allreduce kernels that read and write peer GPUs' fusion buffers directly
// over NVLink 5 / NVSwitch, with the averaging scale and the SGD update fused in.
//
// Replaces the host-driven algorithms of reference src/allreduce.py (a Python isend/recv ring
// over gloo on CPU tensors, :45-98; a broadcast-emulated NCCL ring, :100-170; a central
// gather/scatter, :9-43) and the intra-node host sum of src/reducers.py:59-62.  Behind NVSwitch
// every peer is one hop away at full bandwidth, so there is no ring: every GPU talks to every
// other GPU at once.
//
//   two-shot  rank r pulls slice r of the buffer from all N peers (16-byte loads on peer
//             pointers), sums in fp32 in rank order, scales, and pushes the result into slice r
//             of all N buffers (16-byte stores on peer pointers).  In place; bytes crossing each
//             GPU's links, per direction: 2 (N-1)/N S.  No barrier is needed between the two
//             phases because element i is read and written by exactly one thread in the system.
//   one-shot  every rank pulls the whole buffer from all peers and reduces locally: (N-1) S
//             inbound, one barrier; wins when the buffer is small and latency dominates.
//   nvls      multimem.ld_reduce pulls slice r already summed by the switch, multimem.st
//             broadcasts the result through the switch: about S/N + S per direction.
//
// Device-side barrier: rank r's CTA b writes an epoch number into slot [b][r] of every peer's
// signal pad (st.release.sys after a CTA barrier) and spins (ld.acquire.sys) on its own pad's
// slots [b][*].  Epochs only grow, the comparison is wrap-safe, a rank can be at most one
// barrier ahead of a peer, so one slot per (CTA, peer) is enough and nothing is ever reset.
// The epoch counter lives in device memory, so launches can be captured in a CUDA graph.
// A CTA only ever waits for the *same-numbered* CTA of its peers.  The work is split so that
// this is sufficient: CTA b owns "column b" -- tiles b, b+G, b+2G... of every slice -- both
// when it reduces slice r and when it runs the epilogue over all slices, so everything a CTA
// consumes after a barrier was produced by the CTAs it has just synchronised with.  The waits
// are bounded (dlb_set_wait_timeout_ms): a missing peer sets the error word instead of hanging.
//
// Every CTA of every rank must be resident while it waits, so grids are small (<= 64 CTAs) and
// all ranks must issue the same sequence of launches with the same grid.
//
// Fused epilogue: with `sgd` set, the reduced gradient is consumed by the SGD-momentum update
// in the same kernel -- straight from registers in the one-shot variant.
#include <stdlib.h>

#include <atomic>

#include "sgd_math.cuh"

namespace {

struct CommArgs {
    void* buf[DLB_MAX_RANKS];
    uint32_t* sig[DLB_MAX_RANKS];
    void* out[DLB_MAX_RANKS];  // one-shot result (only [0] outside emulation)
    dlb_fused_sgd_t sgd[DLB_MAX_RANKS];
    uint32_t* state;  // epoch[DLB_MAX_BLOCKS] + error word (emulation: one block of DLB_STATE_WORDS per rank)
    void* mc;         // multicast mapping (nvls)
    int64_t nvec;     // 16-byte vectors in the buffer
    uint64_t timeout_ns;
    float scale;
    int rank, nranks, end_barrier, has_sgd;
};

// 10 minutes, the order of NCCL's watchdog: long enough for a rank that is still loading data or
// timing cuDNN algorithms, short enough that a dead peer ends the job.  $DLB200_WAIT_TIMEOUT_MS overrides.
uint64_t g_timeout_ns = 600000ull * 1000000ull;
bool g_timeout_env_read = false;

uint64_t wait_timeout_ns() {
    if (!g_timeout_env_read) {
        g_timeout_env_read = true;
        if (const char* e = getenv("DLB200_WAIT_TIMEOUT_MS")) {
            const long long ms = atoll(e);
            g_timeout_ns = ms <= 0 ? 0ull : (uint64_t)ms * 1000000ull;
        }
    }
    return g_timeout_ns;
}

template <int NR>
__device__ __forceinline__ int ranks_of(const CommArgs& a) {
    return NR > 0 ? NR : a.nranks;
}

// ---- barrier ------------------------------------------------------------------------------------
template <int NR>
__device__ __forceinline__ void block_barrier(const CommArgs& a, int rank, uint32_t epoch, uint32_t* err) {
    __syncthreads();
    if ((int)threadIdx.x < ranks_of<NR>(a)) {
        const int peer = threadIdx.x;
        // st.release.sys orders the whole CTA's earlier (peer) writes -- made visible to this thread
        // by the __syncthreads above -- before the flag; no separate system fence is needed
        st_flag_release_sys(a.sig[peer] + blockIdx.x * DLB_MAX_RANKS + rank, epoch);
        const uint32_t* mine = a.sig[rank] + blockIdx.x * DLB_MAX_RANKS + peer;
        uint64_t t0 = 0;
        uint32_t spins = 0;
        while ((int32_t)(ld_flag_acquire_sys(mine) - epoch) < 0) {
            if (a.timeout_ns && (++spins & 0x3ff) == 0) {
                const uint64_t now = global_timer_ns();
                if (t0 == 0) {
                    t0 = now;
                } else if (now - t0 > a.timeout_ns) {
                    atomicExch(err, 1u);  // reported by dlb_comm_check; results of this launch are garbage
                    break;
                }
            }
        }
    }
    __syncthreads();
}

template <int NR>
struct Unroll {
    // ~16 peer loads of 16 bytes in flight per thread
    static constexpr int U = NR == 0 ? 2 : (16 / NR > 0 ? 16 / NR : 1);
    static constexpr int R = NR == 0 ? DLB_MAX_RANKS : NR;
};

template <typename T>
__device__ __forceinline__ void accumulate(float* acc, const vec16& v) {
    float f[VecOf<T>::N];
    VecOf<T>::to_float(v, f);
#pragma unroll
    for (int i = 0; i < VecOf<T>::N; ++i) acc[i] += f[i];
}

// The epilogue of two-shot / nvls: SGD over this CTA's column of every slice of the local buffer.
template <typename T, int NR>
__device__ __forceinline__ void sgd_epilogue(const CommArgs& a, const dlb_fused_sgd_t& s, const char* local,
                                             int64_t slice, int64_t tile) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = Unroll<NR>::U;
    const int nr = ranks_of<NR>(a);
    for (int sl = 0; sl < nr; ++sl) {
        const int64_t sbegin = sl * slice, send = min(sbegin + slice, a.nvec);
        for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t j = sbegin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
                if (j < send) {
                    float g[N];
                    VecOf<T>::to_float(ld_vec(local + j * 16), g);
                    sgd_apply_vec<N>(s, (size_t)j * N, g);
                }
            }
        }
    }
}

// ---- two-shot -------------------------------------------------------------------------------------
template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) twoshot_kernel(const __grid_constant__ CommArgs a) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = Unroll<NR>::U;
    constexpr int R = Unroll<NR>::R;
    const int nr = ranks_of<NR>(a);
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    uint32_t* state = a.state + (EMU ? rank * DLB_STATE_WORDS : 0);
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    block_barrier<NR>(a, rank, ++epoch, err);  // every rank's buffer holds its packed gradients

    const int64_t slice = (a.nvec + nr - 1) / nr;
    const int64_t begin = rank * slice, end = min(begin + slice, a.nvec);
    const int64_t tile = (int64_t)blockDim.x * U;
    for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
        vec16 v[U][R];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) {
#pragma unroll
                for (int p = 0; p < R; ++p)
                    if (p < nr) v[u][p] = ld_vec_stream(reinterpret_cast<const char*>(a.buf[p]) + j * 16);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) {
                float acc[N];
                VecOf<T>::to_float(v[u][0], acc);
#pragma unroll
                for (int p = 1; p < R; ++p)
                    if (p < nr) accumulate<T>(acc, v[u][p]);  // fixed rank order: deterministic
#pragma unroll
                for (int i = 0; i < N; ++i) acc[i] *= a.scale;
                const vec16 r = VecOf<T>::from_float(acc);
#pragma unroll
                for (int p = 0; p < R; ++p)
                    if (p < nr) st_vec(reinterpret_cast<char*>(a.buf[p]) + j * 16, r);
            }
        }
    }

    block_barrier<NR>(a, rank, ++epoch, err);  // all slices have landed in every buffer

    if (a.has_sgd)
        sgd_epilogue<T, NR>(a, a.sgd[EMU ? rank : 0], reinterpret_cast<const char*>(a.buf[rank]), slice, tile);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
}

// ---- one-shot -------------------------------------------------------------------------------------
template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) oneshot_kernel(const __grid_constant__ CommArgs a) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = Unroll<NR>::U;
    constexpr int R = Unroll<NR>::R;
    const int nr = ranks_of<NR>(a);
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    uint32_t* state = a.state + (EMU ? rank * DLB_STATE_WORDS : 0);
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];
    char* out = reinterpret_cast<char*>(a.out[EMU ? rank : 0]);
    const dlb_fused_sgd_t& sgd = a.sgd[EMU ? rank : 0];

    block_barrier<NR>(a, rank, ++epoch, err);

    const int64_t tile = (int64_t)blockDim.x * U;
    for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < a.nvec; t0 += (int64_t)gridDim.x * tile) {
        vec16 v[U][R];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < a.nvec) {
#pragma unroll
                for (int p = 0; p < R; ++p)
                    if (p < nr) v[u][p] = ld_vec_stream(reinterpret_cast<const char*>(a.buf[p]) + j * 16);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < a.nvec) {
                float acc[N];
                VecOf<T>::to_float(v[u][0], acc);
#pragma unroll
                for (int p = 1; p < R; ++p)
                    if (p < nr) accumulate<T>(acc, v[u][p]);
#pragma unroll
                for (int i = 0; i < N; ++i) acc[i] *= a.scale;
                if (out) st_vec(out + j * 16, VecOf<T>::from_float(acc));
                if (a.has_sgd) {
                    if (sizeof(T) == 2) {
                        // the update sees the gradient exactly as a later reader of `out` would
                        VecOf<T>::to_float(VecOf<T>::from_float(acc), acc);
                    }
                    sgd_apply_vec<N>(sgd, (size_t)j * N, acc);  // gradient never leaves registers
                }
            }
        }
    }
    if (a.end_barrier) block_barrier<NR>(a, rank, ++epoch, err);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
}

// ---- nvls (multimem) ------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void multimem_ld_reduce(const void* mc, float* acc);
template <typename T>
__device__ __forceinline__ void multimem_st(void* mc, const float* acc);

template <>
__device__ __forceinline__ void multimem_ld_reduce<float>(const void* mc, float* acc) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(acc[0]), "=f"(acc[1]), "=f"(acc[2]), "=f"(acc[3])
                 : "l"(mc)
                 : "memory");
}
template <>
__device__ __forceinline__ void multimem_st<float>(void* mc, const float* acc) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(acc[0]), "f"(acc[1]),
                 "f"(acc[2]), "f"(acc[3])
                 : "memory");
}
template <>
__device__ __forceinline__ void multimem_ld_reduce<__nv_bfloat16>(const void* mc, float* acc) {
    vec16 v;  // the switch accumulates in fp32 and returns bf16x2
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    VecOf<__nv_bfloat16>::to_float(v, acc);
}
template <>
__device__ __forceinline__ void multimem_st<__nv_bfloat16>(void* mc, const float* acc) {
    const vec16 v = VecOf<__nv_bfloat16>::from_float(acc);
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

template <typename T, int NR>
__global__ void __launch_bounds__(512, 1) nvls_kernel(const __grid_constant__ CommArgs a) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = 4;
    const int nr = ranks_of<NR>(a);
    const int rank = a.rank;
    uint32_t* state = a.state;
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    block_barrier<NR>(a, rank, ++epoch, err);

    const int64_t slice = (a.nvec + nr - 1) / nr;
    const int64_t begin = rank * slice, end = min(begin + slice, a.nvec);
    // same column partition as two-shot (tile = blockDim * Unroll<NR>::U) so the epilogue matches
    const int64_t tile = (int64_t)blockDim.x * Unroll<NR>::U;
    char* mc = reinterpret_cast<char*>(a.mc);
    for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
        for (int64_t i0 = 0; i0 < tile; i0 += (int64_t)blockDim.x * U) {
            float acc[U][N];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t off = i0 + (int64_t)u * blockDim.x + threadIdx.x;
                const int64_t j = begin + t0 + off;
                if (off < tile && j < end) multimem_ld_reduce<T>(mc + j * 16, acc[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t off = i0 + (int64_t)u * blockDim.x + threadIdx.x;
                const int64_t j = begin + t0 + off;
                if (off < tile && j < end) {
#pragma unroll
                    for (int i = 0; i < N; ++i) acc[u][i] *= a.scale;
                    multimem_st<T>(mc + j * 16, acc[u]);
                }
            }
        }
    }

    block_barrier<NR>(a, rank, ++epoch, err);

    if (a.has_sgd) sgd_epilogue<T, NR>(a, a.sgd[0], reinterpret_cast<const char*>(a.buf[rank]), slice, tile);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
}

// ---- barrier only ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) barrier_kernel(const __grid_constant__ CommArgs a) {
    uint32_t* state = a.state;
    uint32_t epoch = state[blockIdx.x];
    block_barrier<0>(a, a.rank, ++epoch, state + DLB_MAX_BLOCKS);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
}

// ---- host side ------------------------------------------------------------------------------------
int fill_args(CommArgs& a, void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks, int64_t numel,
              int dtype, float scale, int nblocks, int nthreads, const dlb_fused_sgd_t* sgd, const char* who) {
    DLB_REQUIRE(nranks >= 1 && nranks <= DLB_MAX_RANKS, DLB_E_LIMIT, "%s: nranks %d outside 1..%d", who, nranks, DLB_MAX_RANKS);
    DLB_REQUIRE(rank >= 0 && rank < nranks, DLB_E_ARG, "%s: rank %d of %d", who, rank, nranks);
    DLB_REQUIRE(nblocks >= 1 && nblocks <= DLB_MAX_BLOCKS, DLB_E_LIMIT, "%s: nblocks %d outside 1..%d", who, nblocks, DLB_MAX_BLOCKS);
    DLB_REQUIRE(nthreads >= 32 && nthreads <= 512 && nthreads % 32 == 0, DLB_E_ARG, "%s: nthreads %d (32..512, multiple of 32)", who, nthreads);
    DLB_REQUIRE(dtype == DLB_F32 || dtype == DLB_BF16, DLB_E_ARG, "%s: dtype must be fp32 or bf16", who);
    const int per_vec = 16 / dlb_dtype_size(dtype);
    DLB_REQUIRE(numel >= 0 && numel % per_vec == 0, DLB_E_ARG, "%s: numel %lld is not a multiple of %d (16 bytes)", who, (long long)numel, per_vec);
    DLB_REQUIRE(state != nullptr, DLB_E_ARG, "%s: null state", who);
    for (int r = 0; r < nranks; ++r) {
        DLB_REQUIRE(sigs[r] != nullptr, DLB_E_ARG, "%s: null signal pad for rank %d", who, r);
        DLB_REQUIRE(bufs == nullptr || (bufs[r] != nullptr && (reinterpret_cast<uintptr_t>(bufs[r]) & 15) == 0), DLB_E_ARG,
                    "%s: buffer of rank %d is null or not 16-byte aligned", who, r);
        a.buf[r] = bufs ? bufs[r] : nullptr;
        a.sig[r] = reinterpret_cast<uint32_t*>(sigs[r]);
        a.out[r] = nullptr;
    }
    a.state = state;
    a.mc = nullptr;
    a.nvec = numel / per_vec;
    a.timeout_ns = wait_timeout_ns();
    a.scale = scale;
    a.rank = rank;
    a.nranks = nranks;
    a.end_barrier = 0;
    a.has_sgd = sgd != nullptr;
    if (sgd) {
        DLB_REQUIRE(sgd->param && sgd->momentum_buf, DLB_E_ARG, "%s: fused SGD needs param and momentum", who);
        a.sgd[0] = *sgd;
    }
    return 0;
}

using KernelFn = void (*)(const CommArgs);

int launch(KernelFn k, dim3 grid, int threads, cudaStream_t stream, const CommArgs& a, bool cooperative) {
    void* params[] = {const_cast<CommArgs*>(&a)};
    if (cooperative)
        DLB_CUDA(cudaLaunchCooperativeKernel((const void*)k, grid, dim3(threads), params, 0, stream));
    else
        DLB_CUDA(cudaLaunchKernel((const void*)k, grid, dim3(threads), params, 0, stream));
    return 0;
}

template <typename T, bool EMU>
KernelFn pick_twoshot(int nranks) {
    switch (nranks) {
        case 2: return twoshot_kernel<T, 2, EMU>;
        case 4: return twoshot_kernel<T, 4, EMU>;
        case 8: return twoshot_kernel<T, 8, EMU>;
        default: return twoshot_kernel<T, 0, EMU>;
    }
}
template <typename T, bool EMU>
KernelFn pick_oneshot(int nranks) {
    switch (nranks) {
        case 2: return oneshot_kernel<T, 2, EMU>;
        case 4: return oneshot_kernel<T, 4, EMU>;
        case 8: return oneshot_kernel<T, 8, EMU>;
        default: return oneshot_kernel<T, 0, EMU>;
    }
}
template <typename T>
KernelFn pick_nvls(int nranks) {
    switch (nranks) {
        case 2: return nvls_kernel<T, 2>;
        case 4: return nvls_kernel<T, 4>;
        case 8: return nvls_kernel<T, 8>;
        default: return nvls_kernel<T, 0>;
    }
}

}  // namespace

extern "C" int dlb_set_wait_timeout_ms(int ms) {
    g_timeout_env_read = true;      // an explicit call wins over the environment
    g_timeout_ns = ms <= 0 ? 0ull : (uint64_t)ms * 1000000ull;
    return 0;
}

extern "C" int dlb_comm_check(uint32_t* state, void* stream) {
    uint32_t err = 0;
    DLB_CUDA(cudaMemcpyAsync(&err, state + DLB_MAX_BLOCKS, sizeof(err), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    DLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (err) {
        dlb_set_error("a device-side barrier timed out: a peer GPU never arrived (launch order or grid mismatch, or a dead rank)");
        return DLB_E_TIMEOUT;
    }
    return 0;
}

extern "C" int dlb_allreduce_twoshot(void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks,
                                     int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                     const dlb_fused_sgd_t* sgd, void* stream) {
    CommArgs a;
    DLB_REQUIRE(bufs != nullptr, DLB_E_ARG, "dlb_allreduce_twoshot: null bufs");
    int rc = fill_args(a, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd, "dlb_allreduce_twoshot");
    if (rc || a.nvec == 0) return rc;
    KernelFn k = dtype == DLB_F32 ? pick_twoshot<float, false>(nranks) : pick_twoshot<__nv_bfloat16, false>(nranks);
    return launch(k, dim3(nblocks), nthreads, (cudaStream_t)stream, a, false);
}

extern "C" int dlb_allreduce_oneshot(void* const* bufs, void* out, void* const* sigs, uint32_t* state, int rank,
                                     int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                     int end_barrier, const dlb_fused_sgd_t* sgd, void* stream) {
    CommArgs a;
    DLB_REQUIRE(bufs != nullptr, DLB_E_ARG, "dlb_allreduce_oneshot: null bufs");
    int rc = fill_args(a, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd, "dlb_allreduce_oneshot");
    if (rc || a.nvec == 0) return rc;
    DLB_REQUIRE(out != nullptr || sgd != nullptr, DLB_E_ARG, "dlb_allreduce_oneshot: neither an output buffer nor a fused update");
    DLB_REQUIRE(out != bufs[rank], DLB_E_ARG, "dlb_allreduce_oneshot: out may not alias the input buffer (peers are reading it)");
    DLB_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, DLB_E_ARG, "dlb_allreduce_oneshot: out is not 16-byte aligned");
    a.out[0] = out;
    a.end_barrier = end_barrier;
    KernelFn k = dtype == DLB_F32 ? pick_oneshot<float, false>(nranks) : pick_oneshot<__nv_bfloat16, false>(nranks);
    return launch(k, dim3(nblocks), nthreads, (cudaStream_t)stream, a, false);
}

extern "C" int dlb_allreduce_nvls(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank,
                                  int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                  const dlb_fused_sgd_t* sgd, void* stream) {
    CommArgs a;
    DLB_REQUIRE(bufs != nullptr && mc_buf != nullptr, DLB_E_ARG, "dlb_allreduce_nvls: null buffer");
    DLB_REQUIRE((reinterpret_cast<uintptr_t>(mc_buf) & 15) == 0, DLB_E_ARG, "dlb_allreduce_nvls: multicast pointer not 16-byte aligned");
    int rc = fill_args(a, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd, "dlb_allreduce_nvls");
    if (rc || a.nvec == 0) return rc;
    a.mc = mc_buf;
    KernelFn k = dtype == DLB_F32 ? pick_nvls<float>(nranks) : pick_nvls<__nv_bfloat16>(nranks);
    return launch(k, dim3(nblocks), nthreads, (cudaStream_t)stream, a, false);
}

extern "C" int dlb_barrier(void* const* sigs, uint32_t* state, int rank, int nranks, int nblocks, void* stream) {
    CommArgs a;
    int rc = fill_args(a, nullptr, sigs, state, rank, nranks, 0, DLB_F32, 1.f, nblocks, 32, nullptr, "dlb_barrier");
    if (rc) return rc;
    return launch(barrier_kernel, dim3(nblocks), 32, (cudaStream_t)stream, a, false);
}

// One-shot outputs in the emulation follow dlb_allreduce_oneshot's rules: 16-byte stores, and nobody
// may write memory that a peer is still reading or writing.
static int check_emulated_outs(void* const* bufs, void* const* outs, bool has_sgd, int nranks) {
    DLB_REQUIRE(outs != nullptr || has_sgd, DLB_E_ARG, "dlb_allreduce_emulated: one-shot needs outs or sgds");
    for (int r = 0; r < nranks; ++r) {
        void* o = outs ? outs[r] : nullptr;
        if (o == nullptr) {
            DLB_REQUIRE(has_sgd, DLB_E_ARG, "dlb_allreduce_emulated: rank %d has neither an output nor a fused update", r);
            continue;
        }
        DLB_REQUIRE((reinterpret_cast<uintptr_t>(o) & 15) == 0, DLB_E_ARG,
                    "dlb_allreduce_emulated: output of rank %d is not 16-byte aligned", r);
        for (int q = 0; q < nranks; ++q) {
            DLB_REQUIRE(o != bufs[q], DLB_E_ARG, "dlb_allreduce_emulated: output of rank %d aliases the input buffer of rank %d", r, q);
            DLB_REQUIRE(q >= r || o != outs[q], DLB_E_ARG, "dlb_allreduce_emulated: ranks %d and %d share an output buffer", q, r);
        }
    }
    return 0;
}

extern "C" size_t dlb_emulated_flags_bytes(int nranks) {
    return (size_t)nranks * (DLB_SIGNAL_BYTES + 4 * DLB_STATE_WORDS);
}

// Layout of the emulation scratch: nranks signal pads, then nranks state blocks.
extern "C" int dlb_allreduce_emulated(int algo, void* const* bufs, void* const* outs, const dlb_fused_sgd_t* sgds,
                                      void* flags, int nranks, int64_t numel, int dtype, float scale, int nblocks,
                                      int nthreads, void* stream) {
    DLB_REQUIRE(flags != nullptr && bufs != nullptr, DLB_E_ARG, "dlb_allreduce_emulated: null argument");
    DLB_REQUIRE(algo == 0 || algo == 1, DLB_E_ARG, "dlb_allreduce_emulated: algo must be 0 (two-shot) or 1 (one-shot)");
    DLB_REQUIRE(nranks >= 1 && nranks <= DLB_MAX_RANKS, DLB_E_LIMIT, "dlb_allreduce_emulated: nranks %d outside 1..%d", nranks, DLB_MAX_RANKS);
    void* sigs[DLB_MAX_RANKS];
    char* base = reinterpret_cast<char*>(flags);
    for (int r = 0; r < nranks; ++r) sigs[r] = base + (size_t)r * DLB_SIGNAL_BYTES;
    uint32_t* state = reinterpret_cast<uint32_t*>(base + (size_t)nranks * DLB_SIGNAL_BYTES);
    CommArgs a;
    int rc = fill_args(a, bufs, sigs, state, 0, nranks, numel, dtype, scale, nblocks, nthreads, nullptr, "dlb_allreduce_emulated");
    if (rc || a.nvec == 0) return rc;
    if (sgds) {
        a.has_sgd = 1;
        for (int r = 0; r < nranks; ++r) a.sgd[r] = sgds[r];
    }
    if (algo == 1) {
        rc = check_emulated_outs(bufs, outs, sgds != nullptr, nranks);
        if (rc) return rc;
        for (int r = 0; r < nranks; ++r) a.out[r] = outs ? outs[r] : nullptr;
        a.end_barrier = 1;
    }
    // all nranks * nblocks CTAs must be co-resident: that is what the cooperative launch checks
    KernelFn k;
    if (algo == 0)
        k = dtype == DLB_F32 ? pick_twoshot<float, true>(nranks) : pick_twoshot<__nv_bfloat16, true>(nranks);
    else
        k = dtype == DLB_F32 ? pick_oneshot<float, true>(nranks) : pick_oneshot<__nv_bfloat16, true>(nranks);
    return launch(k, dim3(nblocks, nranks), nthreads, (cudaStream_t)stream, a, true);
}

// ---- dlb_allreduce_emulated_static: the emulated allreduce for callers that cannot allocate ---------
//
// dlb_allreduce_emulated needs a scratch buffer for its barrier flags and epoch counters, and
// expects the caller to bring one.  Some callers have no way to get device memory: a plug-in
// running inside a process whose allocator it does not own (a serving runtime that reserved all
// of HBM at start-up and hands none of it to extensions), a self-test that runs late in a job
// when the device is full, or a C host that links this library but nothing that wraps cudaMalloc.
// For them this variant keeps the scratch inside the library: a fixed array in device memory,
// 35 072 bytes, that every GPU which loads the library holds until the process exits.
//
// What the caller has to know:
//   * Limits.  The array is sized for DLB_MAX_RANKS (8) emulated ranks and DLB_MAX_BLOCKS (64)
//     CTAs per rank; anything larger is refused with DLB_E_LIMIT, nothing is launched.
//   * First call.  The first call on each device looks up the array's device address
//     (cudaGetSymbolAddress, with that device current).  Make that first call before starting
//     any stream capture on the device; later calls only enqueue work on `stream`.
//   * Shared state.  There is one array per device, shared by every call and every thread, and it
//     still holds whatever the previous call left there.  Each call therefore clears the part it
//     uses (cudaMemsetAsync on `stream`, ordered before the kernel) so flags and epoch counters
//     start from zero whatever rank count the previous call used.  Two calls whose work overlaps on
//     the same GPU (different streams, or different host threads) would clear and bump each
//     other's flags: that is a race, and the caller must serialise them.
//   * No allocation.  No cudaMalloc / cudaFree / allocating library call is made, here or below.
namespace {

constexpr size_t kStaticFlagsBytes = (size_t)DLB_MAX_RANKS * (DLB_SIGNAL_BYTES + 4 * DLB_STATE_WORDS);
static __device__ __align__(256) unsigned char g_static_flags[kStaticFlagsBytes];

constexpr int kMaxDevices = 64;
// device address of g_static_flags, per device; host-side cache.  Written once per device and read by
// any thread afterwards: an atomic pointer, so two threads making their first call on a device at
// the same moment both look the address up (same answer) instead of racing on a plain flag.
std::atomic<void*> g_static_flags_ptr[kMaxDevices];

// In host code `g_static_flags` names a host-side shadow of the symbol, not device memory: the
// address a kernel can use has to be asked for, once per device, with that device current.
int static_flags_for_current_device(void** out) {
    int dev = -1;
    DLB_CUDA(cudaGetDevice(&dev));
    DLB_REQUIRE(dev >= 0 && dev < kMaxDevices, DLB_E_LIMIT, "dlb_allreduce_emulated_static: device ordinal %d beyond %d", dev, kMaxDevices - 1);
    void* p = g_static_flags_ptr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        DLB_CUDA(cudaGetSymbolAddress(&p, g_static_flags));
        DLB_REQUIRE(p != nullptr, DLB_E_UNSUPPORTED, "dlb_allreduce_emulated_static: no device address for the scratch array");
        g_static_flags_ptr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

}  // namespace

extern "C" size_t dlb_emulated_static_scratch_bytes(void) { return kStaticFlagsBytes; }

extern "C" int dlb_allreduce_emulated_static(int algo, void* const* bufs, void* const* outs, int nranks, int64_t numel,
                                             int dtype, float scale, int nblocks, int nthreads, void* stream) {
    DLB_REQUIRE(bufs != nullptr, DLB_E_ARG, "dlb_allreduce_emulated_static: null bufs");
    DLB_REQUIRE(algo == 0 || algo == 1, DLB_E_ARG, "dlb_allreduce_emulated_static: algo must be 0 (two-shot) or 1 (one-shot)");
    DLB_REQUIRE(nranks >= 1 && nranks <= DLB_MAX_RANKS, DLB_E_LIMIT,
                "dlb_allreduce_emulated_static: %d ranks, the built-in scratch holds at most %d", nranks, DLB_MAX_RANKS);
    DLB_REQUIRE(nblocks >= 1 && nblocks <= DLB_MAX_BLOCKS, DLB_E_LIMIT,
                "dlb_allreduce_emulated_static: %d CTAs per rank, the built-in scratch holds at most %d", nblocks, DLB_MAX_BLOCKS);
    DLB_REQUIRE(algo == 0 || outs != nullptr, DLB_E_ARG, "dlb_allreduce_emulated_static: one-shot needs outs");
    // everything dlb_allreduce_emulated would refuse is refused here first, so a bad call enqueues
    // nothing at all (not even the clearing of the scratch)
    DLB_REQUIRE(nthreads >= 32 && nthreads <= 512 && nthreads % 32 == 0, DLB_E_ARG,
                "dlb_allreduce_emulated_static: nthreads %d (32..512, multiple of 32)", nthreads);
    DLB_REQUIRE(dtype == DLB_F32 || dtype == DLB_BF16, DLB_E_ARG, "dlb_allreduce_emulated_static: dtype must be fp32 or bf16");
    DLB_REQUIRE(numel >= 0 && numel % (16 / dlb_dtype_size(dtype)) == 0, DLB_E_ARG,
                "dlb_allreduce_emulated_static: numel %lld is not a multiple of 16 bytes' worth", (long long)numel);
    for (int r = 0; r < nranks; ++r)
        DLB_REQUIRE(bufs[r] != nullptr && (reinterpret_cast<uintptr_t>(bufs[r]) & 15) == 0, DLB_E_ARG,
                    "dlb_allreduce_emulated_static: buffer of rank %d is null or not 16-byte aligned", r);
    if (algo == 1) {
        int orc = check_emulated_outs(bufs, outs, false, nranks);
        if (orc) return orc;
    }
    if (numel == 0) return 0;
    void* flags = nullptr;
    int rc = static_flags_for_current_device(&flags);
    if (rc) return rc;
    const size_t used = dlb_emulated_flags_bytes(nranks);
    DLB_REQUIRE(used <= kStaticFlagsBytes, DLB_E_LIMIT, "dlb_allreduce_emulated_static: scratch of %zu bytes exceeds the built-in %zu", used, kStaticFlagsBytes);
    // the array keeps the last call's flags and epochs (possibly laid out for another rank count)
    DLB_CUDA(cudaMemsetAsync(flags, 0, used, (cudaStream_t)stream));
    return dlb_allreduce_emulated(algo, bufs, outs, nullptr, flags, nranks, numel, dtype, scale, nblocks, nthreads, stream);
}

// ---- EXPERIMENTAL: nvls with a deeper load pipeline ---------------------------------------------------
//
// Not selected by default and not yet measured on hardware (written after this round's GPU budget
// was spent; `--algo nvls_deep` or DLB200_EXPERIMENTAL=1 opts in).  Motivation, from the 8-GPU sweep:
// nvls_kernel shares its tile (blockDim x 2 vectors for 8 ranks) with the two-shot kernel so both can
// use one epilogue partition, which leaves two multimem.ld_reduce in flight per thread, and the
// tuner prefers 8-16 CTAs at 4-16 MiB: the variant is latency-bound there (29 % / 59 % of the NVLink
// roofline).  Here the tile is blockDim x 8 vectors with all eight switch reductions issued before
// the first multimem.st, and the epilogue walks the same larger tiles, so the same-numbered-CTA
// argument for the barrier still holds.
namespace {

template <typename T, int NR>
__global__ void __launch_bounds__(512, 1) nvls_deep_kernel(const __grid_constant__ CommArgs a) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = 8;
    const int nr = ranks_of<NR>(a);
    const int rank = a.rank;
    uint32_t* state = a.state;
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    block_barrier<NR>(a, rank, ++epoch, err);

    const int64_t slice = (a.nvec + nr - 1) / nr;
    const int64_t begin = rank * slice, end = min(begin + slice, a.nvec);
    const int64_t tile = (int64_t)blockDim.x * U;
    char* mc = reinterpret_cast<char*>(a.mc);
    for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
        float acc[U][N];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) multimem_ld_reduce<T>(mc + j * 16, acc[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) {
#pragma unroll
                for (int i = 0; i < N; ++i) acc[u][i] *= a.scale;
                multimem_st<T>(mc + j * 16, acc[u]);
            }
        }
    }

    block_barrier<NR>(a, rank, ++epoch, err);

    if (a.has_sgd) {
        const char* local = reinterpret_cast<const char*>(a.buf[rank]);
        for (int sl = 0; sl < nr; ++sl) {
            const int64_t sbegin = sl * slice, send = min(sbegin + slice, a.nvec);
            for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
#pragma unroll 2
                for (int u = 0; u < U; ++u) {
                    const int64_t j = sbegin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
                    if (j < send) {
                        float g[N];
                        VecOf<T>::to_float(ld_vec(local + j * 16), g);
                        sgd_apply_vec<N>(a.sgd[0], (size_t)j * N, g);
                    }
                }
            }
        }
    }
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
}

template <typename T>
KernelFn pick_nvls_deep(int nranks) {
    switch (nranks) {
        case 2: return nvls_deep_kernel<T, 2>;
        case 4: return nvls_deep_kernel<T, 4>;
        case 8: return nvls_deep_kernel<T, 8>;
        default: return nvls_deep_kernel<T, 0>;
    }
}

}  // namespace

extern "C" int dlb_allreduce_nvls_deep(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank,
                                       int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                       const dlb_fused_sgd_t* sgd, void* stream) {
    CommArgs a;
    DLB_REQUIRE(bufs != nullptr && mc_buf != nullptr, DLB_E_ARG, "dlb_allreduce_nvls_deep: null buffer");
    DLB_REQUIRE((reinterpret_cast<uintptr_t>(mc_buf) & 15) == 0, DLB_E_ARG, "dlb_allreduce_nvls_deep: multicast pointer not 16-byte aligned");
    int rc = fill_args(a, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd, "dlb_allreduce_nvls_deep");
    if (rc || a.nvec == 0) return rc;
    a.mc = mc_buf;
    KernelFn k = dtype == DLB_F32 ? pick_nvls_deep<float>(nranks) : pick_nvls_deep<__nv_bfloat16>(nranks);
    return launch(k, dim3(nblocks), nthreads, (cudaStream_t)stream, a, false);
}
