// allreduce.cu -- K-ar: allreduce kernels that read and write peer GPUs' fusion buffers directly
// over NVLink 5 / NVSwitch, with the gradient pack, the averaging scale and the SGD update fused in.
//
// Replaces the host-driven algorithms of reference src/allreduce.py (a Python isend/recv ring
// over gloo on CPU tensors, :45-98; a broadcast-emulated NCCL ring, :100-170; a central
// gather/scatter, :9-43), the intra-node host sum of src/reducers.py:59-62, and -- with the fused
// prologue -- the per-tensor copies of Group.fuse (src/ourdist.py:28-36).  Behind NVSwitch every
// peer is one hop away at full bandwidth, so there is no ring: every GPU talks to every other GPU
// at once.
//
//   two-shot  rank r pulls slice r of the buffer from all N peers (16-byte loads on peer
//             pointers), sums in fp32 in rank order, scales, and pushes the result into slice r
//             of all N buffers (16-byte stores on peer pointers).  In place; bytes crossing each
//             GPU's links, per direction: 2 (N-1)/N S.  No barrier is needed between the two
//             phases because element i is read and written by exactly one thread in the system.
//   one-shot  every rank pulls the whole buffer from all peers and reduces locally: (N-1) S
//             inbound, one barrier; wins when the buffer is small and latency dominates.
//   nvls      multimem.ld_reduce pulls slice r already summed by the switch, multimem.st
//             broadcasts the result through the switch: about S/N + S per direction.  `unroll`
//             (2, 4 or 8) is how many switch reductions a thread has in flight.
//
// One kernel per bucket.  With a tensor table (`ntensors > 0`) the kernel starts by PACKING: CTA b
// gathers, from the gradient tensors' own storage, exactly the tiles of the local buffer that the
// peers' CTAs number b will pull after the barrier ("column b", see below), so the entry barrier
// doubles as "my part of the pack is visible" and no separate pack launch or grid-wide sync is
// needed.  With `sgd` set the kernel ends by consuming the reduced gradient in the SGD-momentum
// update -- straight from registers in the one-shot variant.
//
// Device-side barrier: rank r's CTA b writes an epoch number into slot [b][r] of every peer's
// signal pad (st.release.sys after a CTA barrier) and spins (ld.acquire.sys) on its own pad's
// slots [b][*].  Epochs only grow, the comparison is wrap-safe, a rank can be at most one
// barrier ahead of a peer, so one slot per (CTA, peer) is enough and nothing is ever reset.
// The epoch counter lives in device memory, so launches can be captured in a CUDA graph.
// A CTA only ever waits for the *same-numbered* CTA of its peers.  The work is split so that
// this is sufficient: CTA b owns "column b" -- tiles b, b+G, b+2G... of every slice -- when it
// packs, when it reduces slice r and when it runs the epilogue over all slices, so everything a
// CTA consumes after a barrier was produced by the CTAs it has just synchronised with.  The waits
// are bounded (dlb_set_wait_timeout_ms): a missing peer sets the error word instead of hanging.
//
// Every CTA of every rank must be resident while it waits, so grids are small (<= 64 CTAs) and
// all ranks must issue the same sequence of launches with the same grid, block size and unroll.
//
// `trace` (optional, 8 x uint64 per rank) receives %globaltimer at the phase boundaries of CTA 0:
// start, packed, entry barrier passed, data phase done, exit barrier passed, epilogue done.
#include <stdlib.h>

#include <atomic>

#include "bulk_copy.cuh"
#include "sgd_math.cuh"
#include "tensor_table.cuh"

namespace {

struct CommArgs {
    void* buf[DLB_MAX_RANKS];
    uint32_t* sig[DLB_MAX_RANKS];
    void* out[DLB_MAX_RANKS];  // one-shot result (only [0] outside emulation)
    dlb_fused_sgd_t sgd[DLB_MAX_RANKS];
    uint32_t* state;  // epoch[DLB_MAX_BLOCKS] + error word (emulation: one block of DLB_STATE_WORDS per rank)
    void* mc;         // multicast mapping (nvls)
    unsigned long long* trace;
    int64_t nvec;  // 16-byte vectors in the buffer
    uint64_t timeout_ns;
    float scale, pack_scale;
    int rank, nranks, end_barrier, has_sgd, root;
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

__device__ __forceinline__ void trace_mark(const CommArgs& a, int rank_slot, int i) {
    if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[rank_slot * 8 + i] = global_timer_ns();
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

// ---- column walk ----------------------------------------------------------------------------------
// CTA b's column: tiles b, b+G, ... of each of the `nr` slices.  `q` enumerates them slice-major.
struct Column {
    int64_t slice, tile, nvec;
    int per_slice;  // tiles of one slice that this CTA owns
    int count;      // per_slice * nr
    __device__ __forceinline__ Column(int64_t nvec_, int nr, int64_t tile_) {
        nvec = nvec_;
        tile = tile_;
        slice = (nvec + nr - 1) / nr;
        const int64_t ts = (slice + tile - 1) / tile;  // tiles per slice
        per_slice = (int)blockIdx.x < ts ? (int)((ts - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
        count = per_slice * nr;
    }
    // vector index of (tile q, position u*blockDim + tid), or -1 past the slice / the buffer
    __device__ __forceinline__ int64_t vec_of(int q, int u) const {
        const int sl = q / per_slice, i = q - sl * per_slice;
        const int64_t within = ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * tile + (int64_t)u * blockDim.x + threadIdx.x;
        const int64_t j = sl * slice + within;
        return (within < slice && j < nvec) ? j : -1;
    }
};

// The prologue: gather this CTA's column of the local buffer from the gradient tensors.
// PU tiles x U vectors are in flight per thread.
template <typename T, int U>
__device__ __forceinline__ void pack_columns(const CommArgs& a, const TableSmem& s, int ntensors, char* local, int nr) {
    constexpr int N = VecOf<T>::N;
    constexpr int PU = U >= 8 ? 1 : 8 / U;
    const Column col(a.nvec, nr, (int64_t)blockDim.x * U);
    Owner own;
    for (int q0 = 0; q0 < col.count; q0 += PU) {
        Gathered g[PU][U];
        int64_t j[PU][U];
#pragma unroll
        for (int p = 0; p < PU; ++p) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                j[p][u] = q0 + p < col.count ? col.vec_of(q0 + p, u) : -1;
                g[p][u].kind = -1;
                if (j[p][u] >= 0) gather_issue<T>(s, ntensors, (uint32_t)j[p][u], own, g[p][u]);
            }
        }
#pragma unroll
        for (int p = 0; p < PU; ++p) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float f[N];
                if (j[p][u] >= 0 && gather_finish<T>(g[p][u], f)) {
#pragma unroll
                    for (int i = 0; i < N; ++i) f[i] *= a.pack_scale;
                    st_vec(local + j[p][u] * 16, VecOf<T>::from_float(f));
                }
            }
        }
    }
}

// The epilogue of two-shot / nvls: SGD over this CTA's column of every slice of the local buffer.
template <typename T, int U>
__device__ __forceinline__ void sgd_epilogue(const CommArgs& a, const dlb_fused_sgd_t& s, const char* local, int nr) {
    constexpr int N = VecOf<T>::N;
    const Column col(a.nvec, nr, (int64_t)blockDim.x * U);
    for (int q = 0; q < col.count; ++q) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = col.vec_of(q, u);
            if (j >= 0) {
                float g[N];
                VecOf<T>::to_float(ld_vec(local + j * 16), g);
                sgd_apply_vec<N>(s, (size_t)j * N, g);
            }
        }
    }
}

// ---- two-shot -------------------------------------------------------------------------------------
template <typename T, int NR, bool EMU, bool PACK>
__device__ __forceinline__ void twoshot_body(const CommArgs& a, const TensorTable* tab, TableSmem* ts) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = Unroll<NR>::U;
    constexpr int R = Unroll<NR>::R;
    const int nr = ranks_of<NR>(a);
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int slot = EMU ? rank : 0;
    uint32_t* state = a.state + (EMU ? rank * DLB_STATE_WORDS : 0);
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    trace_mark(a, slot, 0);
    if (PACK) {
        table_stage(*ts, *tab, rank);
        pack_columns<T, U>(a, *ts, tab->ntensors, reinterpret_cast<char*>(a.buf[rank]), nr);
    }
    trace_mark(a, slot, 1);
    block_barrier<NR>(a, rank, ++epoch, err);  // every rank's buffer holds its packed gradients
    trace_mark(a, slot, 2);

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
    trace_mark(a, slot, 3);

    block_barrier<NR>(a, rank, ++epoch, err);  // all slices have landed in every buffer
    trace_mark(a, slot, 4);

    if (a.has_sgd) sgd_epilogue<T, U>(a, a.sgd[slot], reinterpret_cast<const char*>(a.buf[rank]), nr);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
    trace_mark(a, slot, 5);
}

template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) twoshot_kernel(const __grid_constant__ CommArgs a) {
    twoshot_body<T, NR, EMU, false>(a, nullptr, nullptr);
}
template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) twoshot_pack_kernel(const __grid_constant__ CommArgs a,
                                                              const __grid_constant__ TensorTable tab) {
    __shared__ TableSmem ts;
    twoshot_body<T, NR, EMU, true>(a, &tab, &ts);
}

// ---- two-shot through the bulk-copy engine (cp.async.bulk, SASS UBLKCP) ------------------------------
// Same algorithm, same column partition, same result as twoshot_body; what changes is who moves the
// bytes.  One thread asks the copy engine for this CTA's tile of slice `rank` from every peer
// (global -> shared, completion counted on an mbarrier, a three-stage ring keeps two tiles ahead in
// flight), all threads sum the NR staged copies out of shared memory in rank order, and the reduced
// tile leaves as NR bulk stores shared -> peer global.  The SM issues one instruction per peer per
// 8 KB instead of one per 16 bytes, and no register holds data in flight.
constexpr int kBulkArTileVec = 512;  // 16-byte vectors per tile = blockDim.x (one vector per thread)
constexpr int kBulkArStages = 3;

template <int R>
struct BulkArSmem {
    alignas(128) unsigned char in[kBulkArStages][R][kBulkArTileVec * 16];
    alignas(128) unsigned char out[2][kBulkArTileVec * 16];
    alignas(8) uint64_t full[kBulkArStages];
};

template <typename T, int NR, bool EMU, bool PACK>
__device__ __forceinline__ void twoshot_bulk_body(const CommArgs& a, const TensorTable* tab, TableSmem* ts,
                                                  BulkArSmem<Unroll<NR>::R>& sm) {
    constexpr int N = VecOf<T>::N;
    constexpr int R = Unroll<NR>::R;
    const int nr = ranks_of<NR>(a);
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int slot = EMU ? rank : 0;
    uint32_t* state = a.state + (EMU ? rank * DLB_STATE_WORDS : 0);
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    trace_mark(a, slot, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kBulkArStages; ++i) mbar_init(&sm.full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (PACK) {
        table_stage(*ts, *tab, rank);
        pack_columns<T, 1>(a, *ts, tab->ntensors, reinterpret_cast<char*>(a.buf[rank]), nr);
    }
    trace_mark(a, slot, 1);
    block_barrier<NR>(a, rank, ++epoch, err);  // also makes the mbarrier initialisation visible to the CTA
    trace_mark(a, slot, 2);

    const int64_t slice = (a.nvec + nr - 1) / nr;
    const int64_t begin = rank * slice, end = min(begin + slice, a.nvec);
    const int64_t tile = blockDim.x;  // == kBulkArTileVec (checked by the host)
    // this CTA's tiles of its slice: t0 = (blockIdx.x + i * gridDim.x) * tile, i = 0 .. mine-1
    const int64_t have = end > begin ? end - begin : 0;
    const int64_t ntiles = (have + tile - 1) / tile;
    const int mine = (int64_t)blockIdx.x < ntiles ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
    auto tile_at = [&](int i, int64_t& j0, uint32_t& bytes) {
        j0 = begin + ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * tile;
        bytes = (uint32_t)(min(tile, end - j0) * 16);
    };
    auto issue_loads = [&](int i) {
        int64_t j0;
        uint32_t bytes;
        tile_at(i, j0, bytes);
        const int st = i % kBulkArStages;
#pragma unroll
        for (int p = 0; p < R; ++p)
            if (p < nr) bulk_g2s(sm.in[st][p], reinterpret_cast<const char*>(a.buf[p]) + j0 * 16, bytes, &sm.full[st]);
        mbar_arrive_expect_tx(&sm.full[st], bytes * nr);
    };
    if (threadIdx.x == 0) {
        fence_proxy_async_all();  // the flags were acquired through the generic proxy; the engine reads after that
        for (int i = 0; i < mine && i < kBulkArStages - 1; ++i) issue_loads(i);
    }
    for (int i = 0; i < mine; ++i) {
        const int st = i % kBulkArStages;
        if (threadIdx.x == 0) {
            // stage (i-1) % S was read by every thread before the barrier that ended iteration i-1;
            // out[i % 2] was last sent for tile i-2: at most the newest store group may still read
            if (i + kBulkArStages - 1 < mine) issue_loads(i + kBulkArStages - 1);
            bulk_wait_read<1>();
        }
        __syncthreads();
        mbar_wait(&sm.full[st], (i / kBulkArStages) & 1);
        int64_t j0;
        uint32_t bytes;
        tile_at(i, j0, bytes);
        if (threadIdx.x * 16u < bytes) {
            float acc[N];
            VecOf<T>::to_float(reinterpret_cast<const vec16*>(sm.in[st][0])[threadIdx.x], acc);
#pragma unroll
            for (int p = 1; p < R; ++p)
                if (p < nr) accumulate<T>(acc, reinterpret_cast<const vec16*>(sm.in[st][p])[threadIdx.x]);  // rank order
#pragma unroll
            for (int k = 0; k < N; ++k) acc[k] *= a.scale;
            reinterpret_cast<vec16*>(sm.out[i & 1])[threadIdx.x] = VecOf<T>::from_float(acc);
        }
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int p = 0; p < R; ++p)
                if (p < nr) bulk_s2g(reinterpret_cast<char*>(a.buf[p]) + j0 * 16, sm.out[i & 1], bytes);
            bulk_commit();
        }
    }
    if (threadIdx.x == 0) {
        bulk_wait_all<0>();       // every peer store of this CTA has been performed
        fence_proxy_async_all();  // ... and is ordered before the generic-proxy release of the flags
    }
    trace_mark(a, slot, 3);

    block_barrier<NR>(a, rank, ++epoch, err);  // all slices have landed in every buffer
    trace_mark(a, slot, 4);

    if (a.has_sgd) sgd_epilogue<T, 1>(a, a.sgd[slot], reinterpret_cast<const char*>(a.buf[rank]), nr);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
    trace_mark(a, slot, 5);
}

template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) twoshot_bulk_kernel(const __grid_constant__ CommArgs a) {
    extern __shared__ __align__(128) unsigned char bulk_ar_smem[];
    twoshot_bulk_body<T, NR, EMU, false>(a, nullptr, nullptr, *reinterpret_cast<BulkArSmem<Unroll<NR>::R>*>(bulk_ar_smem));
}
template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) twoshot_bulk_pack_kernel(const __grid_constant__ CommArgs a,
                                                                   const __grid_constant__ TensorTable tab) {
    extern __shared__ __align__(128) unsigned char bulk_ar_smem[];
    __shared__ TableSmem ts;
    twoshot_bulk_body<T, NR, EMU, true>(a, &tab, &ts, *reinterpret_cast<BulkArSmem<Unroll<NR>::R>*>(bulk_ar_smem));
}

// ---- one-shot -------------------------------------------------------------------------------------
template <typename T, int NR, bool EMU, bool PACK>
__device__ __forceinline__ void oneshot_body(const CommArgs& a, const TensorTable* tab, TableSmem* ts) {
    constexpr int N = VecOf<T>::N;
    constexpr int U = Unroll<NR>::U;
    constexpr int R = Unroll<NR>::R;
    const int nr = ranks_of<NR>(a);
    const int rank = EMU ? (int)blockIdx.y : a.rank;
    const int slot = EMU ? rank : 0;
    uint32_t* state = a.state + (EMU ? rank * DLB_STATE_WORDS : 0);
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];
    char* out = reinterpret_cast<char*>(a.out[slot]);
    const dlb_fused_sgd_t& sgd = a.sgd[slot];
    const int64_t tile = (int64_t)blockDim.x * U;

    trace_mark(a, slot, 0);
    if (PACK) {
        // one-shot has no slices: CTA b packs and pulls tiles b, b+G, ... of the whole buffer
        constexpr int PU = U >= 8 ? 1 : 8 / U;
        table_stage(*ts, *tab, rank);
        char* local = reinterpret_cast<char*>(a.buf[rank]);
        Owner own;
        for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < a.nvec; t0 += (int64_t)gridDim.x * tile * PU) {
            Gathered g[PU][U];
#pragma unroll
            for (int p = 0; p < PU; ++p) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int64_t j = t0 + (int64_t)p * gridDim.x * tile + (int64_t)u * blockDim.x + threadIdx.x;
                    g[p][u].kind = -1;
                    if (j < a.nvec) gather_issue<T>(*ts, tab->ntensors, (uint32_t)j, own, g[p][u]);
                }
            }
#pragma unroll
            for (int p = 0; p < PU; ++p) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int64_t j = t0 + (int64_t)p * gridDim.x * tile + (int64_t)u * blockDim.x + threadIdx.x;
                    float f[N];
                    if (j < a.nvec && gather_finish<T>(g[p][u], f)) {
#pragma unroll
                        for (int i = 0; i < N; ++i) f[i] *= a.pack_scale;
                        st_vec(local + j * 16, VecOf<T>::from_float(f));
                    }
                }
            }
        }
    }
    trace_mark(a, slot, 1);
    block_barrier<NR>(a, rank, ++epoch, err);
    trace_mark(a, slot, 2);

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
    trace_mark(a, slot, 3);
    if (a.end_barrier) block_barrier<NR>(a, rank, ++epoch, err);
    trace_mark(a, slot, 4);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
    trace_mark(a, slot, 5);
}

template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) oneshot_kernel(const __grid_constant__ CommArgs a) {
    oneshot_body<T, NR, EMU, false>(a, nullptr, nullptr);
}
template <typename T, int NR, bool EMU>
__global__ void __launch_bounds__(512, 1) oneshot_pack_kernel(const __grid_constant__ CommArgs a,
                                                              const __grid_constant__ TensorTable tab) {
    __shared__ TableSmem ts;
    oneshot_body<T, NR, EMU, true>(a, &tab, &ts);
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
// 16-bit: the switch accumulates in fp32 and returns bf16x2; the raw words are kept until all the
// loads of a tile have been issued so that the conversion does not serialise them
__device__ __forceinline__ vec16 multimem_ld_reduce_bf16_raw(const void* mc) {
    vec16 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
template <>
__device__ __forceinline__ void multimem_st<__nv_bfloat16>(void* mc, const float* acc) {
    const vec16 v = VecOf<__nv_bfloat16>::from_float(acc);
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

// UT switch reductions in flight per thread; the tile (blockDim x UT vectors) is also the unit of
// the pack prologue and of the SGD epilogue, so the same-numbered-CTA argument holds for any UT.
template <typename T, int UT, bool PACK>
__device__ __forceinline__ void nvls_body(const CommArgs& a, const TensorTable* tab, TableSmem* ts) {
    constexpr int N = VecOf<T>::N;
    const int nr = a.nranks;
    const int rank = a.rank;
    uint32_t* state = a.state;
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];

    trace_mark(a, 0, 0);
    if (PACK) {
        table_stage(*ts, *tab, 0);
        pack_columns<T, UT>(a, *ts, tab->ntensors, reinterpret_cast<char*>(a.buf[rank]), nr);
    }
    trace_mark(a, 0, 1);
    block_barrier<0>(a, rank, ++epoch, err);
    trace_mark(a, 0, 2);

    const int64_t slice = (a.nvec + nr - 1) / nr;
    const int64_t begin = rank * slice, end = min(begin + slice, a.nvec);
    const int64_t tile = (int64_t)blockDim.x * UT;
    char* mc = reinterpret_cast<char*>(a.mc);
    for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < slice; t0 += (int64_t)gridDim.x * tile) {
        vec16 raw[UT];
#pragma unroll
        for (int u = 0; u < UT; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) {
                if (sizeof(T) == 4) {
                    float f[4];
                    multimem_ld_reduce<float>(mc + j * 16, f);
                    raw[u] = VecOf<float>::from_float(f);
                } else {
                    raw[u] = multimem_ld_reduce_bf16_raw(mc + j * 16);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < UT; ++u) {
            const int64_t j = begin + t0 + (int64_t)u * blockDim.x + threadIdx.x;
            if (j < end) {
                float acc[N];
                VecOf<T>::to_float(raw[u], acc);
#pragma unroll
                for (int i = 0; i < N; ++i) acc[i] *= a.scale;
                multimem_st<T>(mc + j * 16, acc);
            }
        }
    }
    trace_mark(a, 0, 3);

    block_barrier<0>(a, rank, ++epoch, err);
    trace_mark(a, 0, 4);

    if (a.has_sgd) sgd_epilogue<T, UT>(a, a.sgd[0], reinterpret_cast<const char*>(a.buf[rank]), nr);
    if (threadIdx.x == 0) state[blockIdx.x] = epoch;
    trace_mark(a, 0, 5);
}

template <typename T, int UT>
__global__ void __launch_bounds__(512, 1) nvls_kernel(const __grid_constant__ CommArgs a) {
    nvls_body<T, UT, false>(a, nullptr, nullptr);
}
template <typename T, int UT>
__global__ void __launch_bounds__(512, 1) nvls_pack_kernel(const __grid_constant__ CommArgs a,
                                                           const __grid_constant__ TensorTable tab) {
    __shared__ TableSmem ts;
    nvls_body<T, UT, true>(a, &tab, &ts);
}

// ---- broadcast: every rank copies `root`'s buffer into its own ---------------------------------------
// Used at bring-up to give every replica rank 0's parameters without any other communication
// library (reference: replicas rely on the same seed, src/main.py:31-34).
__global__ void __launch_bounds__(512, 1) broadcast_kernel(const __grid_constant__ CommArgs a) {
    uint32_t* state = a.state;
    uint32_t* err = state + DLB_MAX_BLOCKS;
    uint32_t epoch = state[blockIdx.x];
    block_barrier<0>(a, a.rank, ++epoch, err);  // the root's buffer is filled
    if (a.rank != a.root) {
        const char* src = reinterpret_cast<const char*>(a.buf[a.root]);
        char* dst = reinterpret_cast<char*>(a.buf[a.rank]);
        constexpr int U = 8;
        const int64_t tile = (int64_t)blockDim.x * U;
        for (int64_t t0 = (int64_t)blockIdx.x * tile; t0 < a.nvec; t0 += (int64_t)gridDim.x * tile) {
            vec16 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t j = t0 + (int64_t)u * blockDim.x + threadIdx.x;
                if (j < a.nvec) v[u] = ld_vec_stream(src + j * 16);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t j = t0 + (int64_t)u * blockDim.x + threadIdx.x;
                if (j < a.nvec) st_vec(dst + j * 16, v[u]);
            }
        }
    }
    block_barrier<0>(a, a.rank, ++epoch, err);  // everyone has its copy: the root may refill its buffer
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
    DLB_REQUIRE(sigs != nullptr, DLB_E_ARG, "%s: null signal pads", who);
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
    a.trace = nullptr;
    a.nvec = numel / per_vec;
    a.timeout_ns = wait_timeout_ns();
    a.scale = scale;
    a.pack_scale = 1.f;
    a.rank = rank;
    a.nranks = nranks;
    a.end_barrier = 0;
    a.root = 0;
    a.has_sgd = sgd != nullptr;
    if (sgd) {
        DLB_REQUIRE(sgd->param && sgd->momentum_buf, DLB_E_ARG, "%s: fused SGD needs param and momentum", who);
        a.sgd[0] = *sgd;
    }
    return 0;
}

int launch(const void* k, dim3 grid, int threads, cudaStream_t stream, const CommArgs& a, const TensorTable* tab,
           bool cooperative, size_t smem = 0) {
    void* params[] = {const_cast<CommArgs*>(&a), const_cast<TensorTable*>(tab)};
    if (smem > 48 * 1024) DLB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (cooperative)
        DLB_CUDA(cudaLaunchCooperativeKernel(k, grid, dim3(threads), params, smem, stream));
    else
        DLB_CUDA(cudaLaunchKernel(k, grid, dim3(threads), params, smem, stream));
    return 0;
}

#define PICK_NR(KERNEL, T, EMU, nranks)                                     \
    ((nranks) == 2   ? (const void*)KERNEL<T, 2, EMU>                       \
     : (nranks) == 4 ? (const void*)KERNEL<T, 4, EMU>                       \
     : (nranks) == 8 ? (const void*)KERNEL<T, 8, EMU>                       \
                     : (const void*)KERNEL<T, 0, EMU>)

template <bool EMU>
const void* pick_twoshot(int dtype, int nranks, bool pack) {
    if (dtype == DLB_F32) return pack ? PICK_NR(twoshot_pack_kernel, float, EMU, nranks) : PICK_NR(twoshot_kernel, float, EMU, nranks);
    return pack ? PICK_NR(twoshot_pack_kernel, __nv_bfloat16, EMU, nranks) : PICK_NR(twoshot_kernel, __nv_bfloat16, EMU, nranks);
}
template <bool EMU>
const void* pick_oneshot(int dtype, int nranks, bool pack) {
    if (dtype == DLB_F32) return pack ? PICK_NR(oneshot_pack_kernel, float, EMU, nranks) : PICK_NR(oneshot_kernel, float, EMU, nranks);
    return pack ? PICK_NR(oneshot_pack_kernel, __nv_bfloat16, EMU, nranks) : PICK_NR(oneshot_kernel, __nv_bfloat16, EMU, nranks);
}

template <bool EMU>
const void* pick_twoshot_bulk(int dtype, int nranks, bool pack, size_t* smem) {
    *smem = nranks == 2 ? sizeof(BulkArSmem<2>) : nranks == 4 ? sizeof(BulkArSmem<4>) : sizeof(BulkArSmem<DLB_MAX_RANKS>);
    if (dtype == DLB_F32)
        return pack ? PICK_NR(twoshot_bulk_pack_kernel, float, EMU, nranks) : PICK_NR(twoshot_bulk_kernel, float, EMU, nranks);
    return pack ? PICK_NR(twoshot_bulk_pack_kernel, __nv_bfloat16, EMU, nranks) : PICK_NR(twoshot_bulk_kernel, __nv_bfloat16, EMU, nranks);
}

#define PICK_UT(KERNEL, T, ut) ((ut) == 2 ? (const void*)KERNEL<T, 2> : (ut) == 4 ? (const void*)KERNEL<T, 4> : (const void*)KERNEL<T, 8>)

const void* pick_nvls(int dtype, int unroll, bool pack) {
    if (dtype == DLB_F32) return pack ? PICK_UT(nvls_pack_kernel, float, unroll) : PICK_UT(nvls_kernel, float, unroll);
    return pack ? PICK_UT(nvls_pack_kernel, __nv_bfloat16, unroll) : PICK_UT(nvls_kernel, __nv_bfloat16, unroll);
}

// Fill the kernel's tensor table for the fused pack prologue.  Returns 0 and has_pack = false when
// there is nothing to pack.
int fill_pack(TensorTable& tab, bool* has_pack, int dtype, int64_t numel, void* const* tensors, const int* tensor_dtypes,
              const int64_t* offsets, const int64_t* numels, int ntensors, const char* who) {
    *has_pack = false;
    if (ntensors <= 0) return 0;
    DLB_REQUIRE(ntensors <= DLB_MAX_TENSORS, DLB_E_LIMIT, "%s: %d tensors in the fused pack, at most %d (pack separately)", who,
                ntensors, DLB_MAX_TENSORS);
    DLB_REQUIRE(tensors && tensor_dtypes && offsets && numels, DLB_E_ARG, "%s: null tensor table", who);
    bool aligned = false;
    uint32_t first = 0, end = 0;
    int rc = dlb_fill_table(tab, dtype, tensors, tensor_dtypes, offsets, numels, 0, ntensors, &aligned, &first, &end);
    if (rc) return rc;
    if (tab.ntensors == 0) return 0;
    DLB_REQUIRE(aligned, DLB_E_ARG, "%s: the fused pack needs ascending, non-overlapping offsets on 16-byte boundaries of the buffer", who);
    DLB_REQUIRE((int64_t)end <= numel, DLB_E_ARG, "%s: tensor table reaches element %u, beyond the buffer's %lld", who, end, (long long)numel);
    *has_pack = true;
    return 0;
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

// ---- the general entry point ----------------------------------------------------------------------------
extern "C" int dlb_allreduce(const dlb_allreduce_desc_t* d, void* stream) {
    DLB_REQUIRE(d != nullptr, DLB_E_ARG, "dlb_allreduce: null descriptor");
    const char* who = d->algo == DLB_ALGO_TWOSHOT ? "dlb_allreduce(twoshot)" : d->algo == DLB_ALGO_ONESHOT ? "dlb_allreduce(oneshot)"
                      : d->algo == DLB_ALGO_NVLS  ? "dlb_allreduce(nvls)" : "dlb_allreduce(twoshot_bulk)";
    DLB_REQUIRE(d->algo >= DLB_ALGO_TWOSHOT && d->algo <= DLB_ALGO_TWOSHOT_BULK, DLB_E_ARG, "dlb_allreduce: unknown algo %d", d->algo);
    DLB_REQUIRE(d->bufs != nullptr, DLB_E_ARG, "%s: null bufs", who);
    CommArgs a;
    int rc = fill_args(a, d->bufs, d->sigs, d->state, d->rank, d->nranks, d->numel, d->dtype, d->scale, d->nblocks, d->nthreads, d->sgd, who);
    if (rc || a.nvec == 0) return rc;
    const int unroll = d->unroll == 0 ? 4 : d->unroll;
    DLB_REQUIRE(d->algo != DLB_ALGO_NVLS || unroll == 2 || unroll == 4 || unroll == 8, DLB_E_ARG, "%s: unroll %d (2, 4 or 8)", who, unroll);
    TensorTable tab;
    bool pack = false;
    rc = fill_pack(tab, &pack, d->dtype, d->numel, d->tensors, d->tensor_dtypes, d->offsets, d->numels, d->ntensors, who);
    if (rc) return rc;
    a.pack_scale = d->pack_scale == 0.f ? 1.f : d->pack_scale;
    a.trace = d->trace;
    const void* k = nullptr;
    size_t smem = 0;
    if (d->algo == DLB_ALGO_TWOSHOT) {
        k = pick_twoshot<false>(d->dtype, d->nranks, pack);
    } else if (d->algo == DLB_ALGO_TWOSHOT_BULK) {
        DLB_REQUIRE(d->nthreads == kBulkArTileVec, DLB_E_ARG, "%s: needs %d threads per CTA (one 16-byte vector of the tile each)", who, kBulkArTileVec);
        k = pick_twoshot_bulk<false>(d->dtype, d->nranks, pack, &smem);
    } else if (d->algo == DLB_ALGO_ONESHOT) {
        DLB_REQUIRE(d->out != nullptr || d->sgd != nullptr, DLB_E_ARG, "%s: neither an output buffer nor a fused update", who);
        DLB_REQUIRE(d->out != d->bufs[d->rank], DLB_E_ARG, "%s: out may not alias the input buffer (peers are reading it)", who);
        DLB_REQUIRE((reinterpret_cast<uintptr_t>(d->out) & 15) == 0, DLB_E_ARG, "%s: out is not 16-byte aligned", who);
        a.out[0] = d->out;
        a.end_barrier = d->end_barrier;
        k = pick_oneshot<false>(d->dtype, d->nranks, pack);
    } else {
        DLB_REQUIRE(d->mc_buf != nullptr && (reinterpret_cast<uintptr_t>(d->mc_buf) & 15) == 0, DLB_E_ARG,
                    "%s: multicast pointer null or not 16-byte aligned", who);
        a.mc = d->mc_buf;
        k = pick_nvls(d->dtype, unroll, pack);
    }
    return launch(k, dim3(d->nblocks), d->nthreads, (cudaStream_t)stream, a, pack ? &tab : nullptr, false, smem);
}

// ---- the original entry points: thin wrappers ------------------------------------------------------------
static void base_desc(dlb_allreduce_desc_t& d, int algo, void* const* bufs, void* const* sigs, uint32_t* state, int rank,
                      int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads, const dlb_fused_sgd_t* sgd) {
    d = dlb_allreduce_desc_t{};
    d.algo = algo;
    d.bufs = bufs;
    d.sigs = sigs;
    d.state = state;
    d.rank = rank;
    d.nranks = nranks;
    d.numel = numel;
    d.dtype = dtype;
    d.scale = scale;
    d.nblocks = nblocks;
    d.nthreads = nthreads;
    d.sgd = sgd;
    d.pack_scale = 1.f;
}

extern "C" int dlb_allreduce_twoshot(void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks,
                                     int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                     const dlb_fused_sgd_t* sgd, void* stream) {
    dlb_allreduce_desc_t d;
    base_desc(d, DLB_ALGO_TWOSHOT, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd);
    return dlb_allreduce(&d, stream);
}

extern "C" int dlb_allreduce_oneshot(void* const* bufs, void* out, void* const* sigs, uint32_t* state, int rank,
                                     int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                     int end_barrier, const dlb_fused_sgd_t* sgd, void* stream) {
    dlb_allreduce_desc_t d;
    base_desc(d, DLB_ALGO_ONESHOT, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd);
    d.out = out;
    d.end_barrier = end_barrier;
    return dlb_allreduce(&d, stream);
}

extern "C" int dlb_allreduce_nvls(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank,
                                  int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                  const dlb_fused_sgd_t* sgd, void* stream) {
    dlb_allreduce_desc_t d;
    base_desc(d, DLB_ALGO_NVLS, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd);
    d.mc_buf = mc_buf;
    d.unroll = 4;
    return dlb_allreduce(&d, stream);
}

// the "deep" variant of round 1 is the same kernel with eight reductions in flight per thread
extern "C" int dlb_allreduce_nvls_deep(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank,
                                       int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                                       const dlb_fused_sgd_t* sgd, void* stream) {
    dlb_allreduce_desc_t d;
    base_desc(d, DLB_ALGO_NVLS, bufs, sigs, state, rank, nranks, numel, dtype, scale, nblocks, nthreads, sgd);
    d.mc_buf = mc_buf;
    d.unroll = 8;
    return dlb_allreduce(&d, stream);
}

extern "C" int dlb_barrier(void* const* sigs, uint32_t* state, int rank, int nranks, int nblocks, void* stream) {
    CommArgs a;
    int rc = fill_args(a, nullptr, sigs, state, rank, nranks, 0, DLB_F32, 1.f, nblocks, 32, nullptr, "dlb_barrier");
    if (rc) return rc;
    return launch((const void*)barrier_kernel, dim3(nblocks), 32, (cudaStream_t)stream, a, nullptr, false);
}

extern "C" int dlb_broadcast(void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks, int root,
                             int64_t nbytes, int nblocks, int nthreads, void* stream) {
    CommArgs a;
    DLB_REQUIRE(bufs != nullptr, DLB_E_ARG, "dlb_broadcast: null bufs");
    DLB_REQUIRE(root >= 0 && root < nranks, DLB_E_ARG, "dlb_broadcast: root %d of %d", root, nranks);
    DLB_REQUIRE(nbytes >= 0 && nbytes % 16 == 0, DLB_E_ARG, "dlb_broadcast: %lld bytes is not a multiple of 16", (long long)nbytes);
    int rc = fill_args(a, bufs, sigs, state, rank, nranks, nbytes / 4, DLB_F32, 1.f, nblocks, nthreads, nullptr, "dlb_broadcast");
    if (rc || a.nvec == 0) return rc;
    a.root = root;
    return launch((const void*)broadcast_kernel, dim3(nblocks), nthreads, (cudaStream_t)stream, a, nullptr, false);
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
extern "C" int dlb_allreduce_emulated_ex(const dlb_emulated_desc_t* d, void* stream) {
    DLB_REQUIRE(d != nullptr, DLB_E_ARG, "dlb_allreduce_emulated: null descriptor");
    DLB_REQUIRE(d->flags != nullptr && d->bufs != nullptr, DLB_E_ARG, "dlb_allreduce_emulated: null argument");
    DLB_REQUIRE(d->algo == DLB_ALGO_TWOSHOT || d->algo == DLB_ALGO_ONESHOT || d->algo == DLB_ALGO_TWOSHOT_BULK, DLB_E_ARG,
                "dlb_allreduce_emulated: algo must be 0 (two-shot), 1 (one-shot) or 3 (two-shot through the bulk-copy engine)");
    DLB_REQUIRE(d->algo != DLB_ALGO_TWOSHOT_BULK || d->nthreads == kBulkArTileVec, DLB_E_ARG,
                "dlb_allreduce_emulated: the bulk-copy variant needs %d threads per CTA", kBulkArTileVec);
    const int nranks = d->nranks;
    DLB_REQUIRE(nranks >= 1 && nranks <= DLB_MAX_RANKS, DLB_E_LIMIT, "dlb_allreduce_emulated: nranks %d outside 1..%d", nranks, DLB_MAX_RANKS);
    void* sigs[DLB_MAX_RANKS];
    char* base = reinterpret_cast<char*>(d->flags);
    for (int r = 0; r < nranks; ++r) sigs[r] = base + (size_t)r * DLB_SIGNAL_BYTES;
    uint32_t* state = reinterpret_cast<uint32_t*>(base + (size_t)nranks * DLB_SIGNAL_BYTES);
    CommArgs a;
    int rc = fill_args(a, d->bufs, sigs, state, 0, nranks, d->numel, d->dtype, d->scale, d->nblocks, d->nthreads, nullptr, "dlb_allreduce_emulated");
    if (rc || a.nvec == 0) return rc;
    if (d->sgds) {
        a.has_sgd = 1;
        for (int r = 0; r < nranks; ++r) a.sgd[r] = d->sgds[r];
    }
    if (d->algo == DLB_ALGO_ONESHOT) {
        rc = check_emulated_outs(d->bufs, d->outs, d->sgds != nullptr, nranks);
        if (rc) return rc;
        for (int r = 0; r < nranks; ++r) a.out[r] = d->outs ? d->outs[r] : nullptr;
        a.end_barrier = 1;
    }
    TensorTable tab;
    bool pack = false;
    rc = fill_pack(tab, &pack, d->dtype, d->numel, d->tensors, d->tensor_dtypes, d->offsets, d->numels, d->ntensors, "dlb_allreduce_emulated");
    if (rc) return rc;
    tab.emu_stacked = 1;
    a.pack_scale = d->pack_scale == 0.f ? 1.f : d->pack_scale;
    a.trace = d->trace;
    // all nranks * nblocks CTAs must be co-resident: that is what the cooperative launch checks
    size_t smem = 0;
    const void* k = d->algo == DLB_ALGO_TWOSHOT        ? pick_twoshot<true>(d->dtype, nranks, pack)
                    : d->algo == DLB_ALGO_TWOSHOT_BULK ? pick_twoshot_bulk<true>(d->dtype, nranks, pack, &smem)
                                                       : pick_oneshot<true>(d->dtype, nranks, pack);
    return launch(k, dim3(d->nblocks, nranks), d->nthreads, (cudaStream_t)stream, a, pack ? &tab : nullptr, true, smem);
}

extern "C" int dlb_allreduce_emulated(int algo, void* const* bufs, void* const* outs, const dlb_fused_sgd_t* sgds,
                                      void* flags, int nranks, int64_t numel, int dtype, float scale, int nblocks,
                                      int nthreads, void* stream) {
    dlb_emulated_desc_t d = {};
    d.algo = algo;
    d.bufs = bufs;
    d.outs = outs;
    d.sgds = sgds;
    d.flags = flags;
    d.nranks = nranks;
    d.numel = numel;
    d.dtype = dtype;
    d.scale = scale;
    d.nblocks = nblocks;
    d.nthreads = nthreads;
    d.pack_scale = 1.f;
    return dlb_allreduce_emulated_ex(&d, stream);
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
