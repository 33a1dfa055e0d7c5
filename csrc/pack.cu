// pack.cu -- K-pack / K-unpack: multi-tensor gather(+scale+cast) into a flat fusion buffer
// and the reverse scatter.
//
// Replaces reference Group.fuse / Group.unfuse (src/ourdist.py:28-43), which issue one ATen copy
// per tensor (187 copies per step for GoogLeNet, each a blocking D2H when the buffer is on the
// host) followed by a separate full-buffer divide (src/allreduce.py:98).  One launch handles up
// to DLB_MAX_TENSORS tensors; the tensor table travels in the kernel parameters.
//
// Three kernels, chosen by the host (dlb_pack_ex's `variant`, 0 = pick):
//
//   tiles  CTAs walk the FLAT buffer in equal tiles; every thread keeps U 16-byte vectors in
//          flight, finding the owner of each vector in the shared-memory copy of the table
//          (tensor_table.cuh).  Work per CTA is the same whatever the tensor sizes are, so
//          GoogLeNet's 4-byte biases cost nothing extra.  The same code is the allreduce kernels'
//          fused prologue.
//   bulk   (same layouts, tensor dtype == flat dtype)  The Blackwell/Hopper bulk-copy engine moves
//          the data: one thread issues cp.async.bulk global->shared per contiguous piece of a tile
//          (completion on an mbarrier), the CTA scales in shared memory only if scale != 1, and the
//          tile leaves with cp.async.bulk shared->global.  A four-stage ring keeps three tiles of
//          loads in flight per CTA with no registers holding data.  SASS: UBLKCP.
//   chunks (any layout, the round-1 kernel)  One CTA per 8192-element chunk of one tensor; used
//          when offsets are not 16-byte aligned in the flat buffer.
//
// Bandwidth bound: bytes = numel * (src + dst element size).
#include <stdlib.h>

#include <algorithm>
#include <type_traits>

#include "bulk_copy.cuh"
#include "tensor_table.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 8192;  // elements per CTA (chunks kernel)
constexpr int kUnroll = 8;    // 16-byte groups in flight per thread

int sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            sms = 148;
    }
    return sms;
}

// =====================================================================================================
// tiles kernel
// =====================================================================================================
template <typename FlatT, bool PACK, int U>
__global__ void __launch_bounds__(kThreads) pack_tiles_kernel(const __grid_constant__ TensorTable tab,
                                                              FlatT* __restrict__ flat, uint32_t v_begin,
                                                              uint32_t v_end, float scale) {
    constexpr int N = VecOf<FlatT>::N;
    __shared__ TableSmem s;
    table_stage(s, tab, 0);
    const int n = tab.ntensors;
    const uint32_t tile = kThreads * U;
    Owner own;
    for (uint32_t v0 = v_begin + blockIdx.x * tile; v0 < v_end; v0 += gridDim.x * tile) {
        if (PACK) {
            Gathered g[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t v = v0 + u * kThreads + threadIdx.x;
                g[u].kind = -1;
                if (v < v_end) gather_issue<FlatT>(s, n, v, own, g[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t v = v0 + u * kThreads + threadIdx.x;
                float f[N];
                if (gather_finish<FlatT>(g[u], f)) {
#pragma unroll
                    for (int i = 0; i < N; ++i) f[i] *= scale;
                    st_vec(flat + (size_t)v * N, VecOf<FlatT>::from_float(f));
                }
            }
        } else {
            vec16 w[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t v = v0 + u * kThreads + threadIdx.x;
                if (v < v_end) w[u] = ld_vec_stream(flat + (size_t)v * N);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t v = v0 + u * kThreads + threadIdx.x;
                if (v < v_end) {
                    float f[N];
                    VecOf<FlatT>::to_float(w[u], f);
#pragma unroll
                    for (int i = 0; i < N; ++i) f[i] *= scale;
                    scatter_vec<FlatT>(s, n, v, own, f);
                }
            }
        }
    }
}

// =====================================================================================================
// bulk kernel (cp.async.bulk + mbarrier)
// =====================================================================================================
constexpr int kBulkThreads = 128;
constexpr int kBulkTileBytes = 16384;
constexpr int kBulkStages = 4;

struct BulkSmem {
    alignas(128) unsigned char data[kBulkStages][kBulkTileBytes];
    alignas(8) uint64_t full[kBulkStages];
    TableSmem tab;
};

// A piece of one tile: elements [lo, hi) of the flat buffer that belong to tensor t.
struct Piece {
    uint32_t lo, hi;   // flat elements; lo is a vector boundary
    uint32_t nfull;    // elements in whole vectors that the copy engine can move (0: threads move it all)
    char* tensor_at;   // tensor storage at element `lo` (nullptr: no storage)
};

template <typename FlatT>
__device__ __forceinline__ Piece make_piece(const TableSmem& s, int t, uint32_t e0, uint32_t e1) {
    constexpr int N = VecOf<FlatT>::N;
    Piece p;
    const uint32_t off = s.offset[t], end = off + s.numel[t];
    p.lo = max(e0, off);
    p.hi = min(e1, end);
    char* base = reinterpret_cast<char*>(s.tensor[t]);
    p.tensor_at = base ? base + (size_t)(p.lo - off) * sizeof(FlatT) : nullptr;
    const bool engine = base != nullptr && (reinterpret_cast<uintptr_t>(p.tensor_at) & 15) == 0;
    p.nfull = engine ? (p.hi - p.lo) / N * N : 0;
    return p;
}

// Same dtype on both sides.  PACK: tensors -> flat; else flat -> tensors.
template <typename FlatT, bool PACK>
__global__ void __launch_bounds__(kBulkThreads) pack_bulk_kernel(const __grid_constant__ TensorTable tab,
                                                                 FlatT* __restrict__ flat, uint32_t v_begin,
                                                                 uint32_t v_end, float scale) {
    constexpr int N = VecOf<FlatT>::N;
    constexpr uint32_t TILE_V = kBulkTileBytes / 16;  // vectors per tile
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BulkSmem& sm = *reinterpret_cast<BulkSmem*>(smem_raw);
    const int n = tab.ntensors;
    table_stage(sm.tab, tab, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kBulkStages; ++i) mbar_init(&sm.full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t ntiles = (v_end - v_begin + TILE_V - 1) / TILE_V;
    const uint32_t mine = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto tile_range = [&](uint32_t it, uint32_t& e0, uint32_t& e1) {
        const uint32_t v0 = v_begin + (blockIdx.x + it * gridDim.x) * TILE_V;
        e0 = v0 * N;
        e1 = min(v0 + TILE_V, v_end) * N;
    };
    // thread 0: start the loads of this CTA's tile `it` into stage it % kBulkStages
    auto issue_loads = [&](uint32_t it) {
        uint32_t e0, e1;
        tile_range(it, e0, e1);
        const int st = it % kBulkStages;
        unsigned char* dst = sm.data[st];
        uint32_t total = 0;
        if (PACK) {
            for (int t = table_find(sm.tab, n, e0, -1); t < n && sm.tab.offset[t] < e1; ++t) {
                const Piece p = make_piece<FlatT>(sm.tab, t, e0, e1);
                if (p.hi <= p.lo || p.nfull == 0) continue;
                const uint32_t bytes = p.nfull * sizeof(FlatT);
                bulk_g2s(dst + (size_t)(p.lo - e0) * sizeof(FlatT), p.tensor_at, bytes, &sm.full[st]);
                total += bytes;
            }
        } else {
            total = (e1 - e0) * sizeof(FlatT);
            bulk_g2s(dst, flat + e0, total, &sm.full[st]);
        }
        mbar_arrive_expect_tx(&sm.full[st], total);
    };

    constexpr uint32_t kAhead = kBulkStages - 2;  // tiles of loads in flight beyond the current one
    if (threadIdx.x == 0)
        for (uint32_t it = 0; it < mine && it < kAhead; ++it) issue_loads(it);

    for (uint32_t it = 0; it < mine; ++it) {
        const int st = it % kBulkStages;
        if (threadIdx.x == 0 && it + kAhead < mine) {
            // the stage about to be refilled held tile it-2: its stores were committed two iterations
            // ago, so only the most recent store group may still be reading shared memory
            bulk_wait_read<1>();
            issue_loads(it + kAhead);
        }
        mbar_wait(&sm.full[st], (it / kBulkStages) & 1);
        uint32_t e0, e1;
        tile_range(it, e0, e1);
        FlatT* tile_smem = reinterpret_cast<FlatT*>(sm.data[st]);

        // work the copy engine cannot do: zeros for tensors without storage, pieces whose storage is
        // not 16-byte aligned, the partial last vector of a tensor, and the scale
        bool touched = false;
        if (PACK) {
            for (int t = table_find(sm.tab, n, e0, -1); t < n && sm.tab.offset[t] < e1; ++t) {
                const Piece p = make_piece<FlatT>(sm.tab, t, e0, e1);
                if (p.hi <= p.lo) continue;
                const uint32_t hi_pad = min(e1, (p.hi + N - 1) / N * N);  // zero the pad of a partial vector
                if (p.lo + p.nfull < hi_pad) {
                    touched = true;
                    const FlatT* src = reinterpret_cast<const FlatT*>(p.tensor_at);
                    for (uint32_t e = p.lo + p.nfull + threadIdx.x; e < hi_pad; e += kBulkThreads) {
                        const float x = (src != nullptr && e < p.hi) ? VecOf<FlatT>::load1(src + (e - p.lo)) : 0.f;
                        VecOf<FlatT>::store1(tile_smem + (e - e0), x);
                    }
                }
            }
        }
        if (scale != 1.f) {
            touched = true;
            __syncthreads();
            for (uint32_t v = threadIdx.x; v < (e1 - e0) / N; v += kBulkThreads) {
                vec16* q = reinterpret_cast<vec16*>(tile_smem) + v;
                float f[N];
                VecOf<FlatT>::to_float(*q, f);
#pragma unroll
                for (int i = 0; i < N; ++i) f[i] *= scale;
                *q = VecOf<FlatT>::from_float(f);
            }
        }
        if (touched) fence_proxy_async_smem();
        __syncthreads();

        if (PACK) {
            if (threadIdx.x == 0) {
                // one store per maximal run of adjacent pieces (holes in the layout are not written)
                uint32_t run_lo = 0, run_hi = 0;
                for (int t = table_find(sm.tab, n, e0, -1); t < n && sm.tab.offset[t] < e1; ++t) {
                    const Piece p = make_piece<FlatT>(sm.tab, t, e0, e1);
                    if (p.hi <= p.lo) continue;
                    const uint32_t hi_pad = min(e1, (p.hi + N - 1) / N * N);
                    if (run_hi == p.lo && run_hi > run_lo) {
                        run_hi = hi_pad;
                    } else {
                        if (run_hi > run_lo) bulk_s2g(flat + run_lo, tile_smem + (run_lo - e0), (run_hi - run_lo) * sizeof(FlatT));
                        run_lo = p.lo;
                        run_hi = hi_pad;
                    }
                }
                if (run_hi > run_lo) bulk_s2g(flat + run_lo, tile_smem + (run_lo - e0), (run_hi - run_lo) * sizeof(FlatT));
                bulk_commit();
            }
        } else {
            // flat -> tensors: whole vectors through the engine, the rest by threads straight from shared memory
            for (int t = table_find(sm.tab, n, e0, -1); t < n && sm.tab.offset[t] < e1; ++t) {
                const Piece p = make_piece<FlatT>(sm.tab, t, e0, e1);
                if (p.hi <= p.lo || p.tensor_at == nullptr) continue;
                if (threadIdx.x == 0 && p.nfull)
                    bulk_s2g(p.tensor_at, tile_smem + (p.lo - e0), p.nfull * sizeof(FlatT));
                FlatT* dst = reinterpret_cast<FlatT*>(p.tensor_at);
                for (uint32_t e = p.lo + p.nfull + threadIdx.x; e < p.hi; e += kBulkThreads) dst[e - p.lo] = tile_smem[e - e0];
            }
            if (threadIdx.x == 0) bulk_commit();
            // the threads' reads of this stage must be over before thread 0 refills it next iteration
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) bulk_wait_all<0>();
}

// =====================================================================================================
// chunks kernel (any layout)
// =====================================================================================================
struct ChunkTable {
    TensorTable t;
    uint32_t chunk_prefix[DLB_MAX_TENSORS + 1];
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
__global__ void __launch_bounds__(kThreads) pack_chunks_kernel(const __grid_constant__ ChunkTable tab,
                                                               FlatT* __restrict__ flat, float scale) {
    const uint32_t b = blockIdx.x;
    int lo = 0, hi = tab.t.ntensors;  // invariant: chunk_prefix[lo] <= b < chunk_prefix[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tab.chunk_prefix[mid] <= b) lo = mid; else hi = mid;
    }
    const int t = lo;
    const uint32_t start = (b - tab.chunk_prefix[t]) * (uint32_t)kChunk;
    const uint32_t n = tab.t.numel[t];
    const int cnt = (int)min((uint32_t)kChunk, n - start);
    FlatT* dst = flat + tab.t.offset[t];
    void* tensor = tab.t.tensor[t];
    switch (tab.t.dtype[t]) {
        case DLB_F32: move_chunk<FlatT, float, PACK>(tensor, dst, start, cnt, scale); break;
        case DLB_BF16: move_chunk<FlatT, __nv_bfloat16, PACK>(tensor, dst, start, cnt, scale); break;
        default: move_chunk<FlatT, __half, PACK>(tensor, dst, start, cnt, scale); break;
    }
}

// =====================================================================================================
// host
// =====================================================================================================
enum { kAuto = 0, kChunks = 1, kTiles = 2, kBulk = 3 };

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

template <typename FlatT, bool PACK, int U>
void launch_tiles(const TensorTable& tab, FlatT* flat, uint32_t v_begin, uint32_t v_end, float scale, int per_sm,
                  cudaStream_t stream) {
    auto kern = pack_tiles_kernel<FlatT, PACK, U>;
    static int resident = 0;
    if (resident == 0 && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kern, kThreads, 0) != cudaSuccess) resident = 4;
    const uint32_t tiles = (v_end - v_begin + kThreads * U - 1) / (kThreads * U);
    // one CTA per tile: measured faster than a resident-sized persistent grid (the hardware scheduler
    // overlaps the load and store phases of successive CTAs); DLB200_PACK_CTAS_PER_SM caps the grid
    (void)resident;
    const uint32_t grid = per_sm > 0 ? std::min<uint32_t>(tiles, (uint32_t)sm_count() * per_sm) : tiles;
    kern<<<grid, kThreads, 0, stream>>>(tab, flat, v_begin, v_end, scale);
}

template <typename FlatT, bool PACK>
int launch_one(const ChunkTable& ct, int variant, bool aligned, bool same_dtype, uint32_t first, uint32_t end, FlatT* flat,
               float scale, cudaStream_t stream) {
    constexpr int N = VecOf<FlatT>::N;
    const TensorTable& tab = ct.t;
    if (variant == kAuto) variant = aligned ? kTiles : kChunks;
    DLB_REQUIRE(variant == kChunks || aligned, DLB_E_ARG,
                "this pack variant needs ascending, non-overlapping offsets that are multiples of 16 bytes of the flat dtype");
    DLB_REQUIRE(variant != kBulk || same_dtype, DLB_E_ARG, "the bulk-copy pack variant moves bytes: tensor dtype must equal the flat dtype");
    if (variant == kChunks) {
        const uint32_t chunks = ct.chunk_prefix[tab.ntensors];
        if (chunks == 0) return 0;
        pack_chunks_kernel<FlatT, PACK><<<chunks, kThreads, 0, stream>>>(ct, flat, scale);
    } else {
        const uint32_t v_begin = first / N, v_end = (end + N - 1) / N;
        if (v_end <= v_begin) return 0;
        if (variant == kTiles) {
            // persistent grid = exactly what is resident, so there is no second wave
            static const int u_env = env_int("DLB200_PACK_U", 8);
            static const int per_sm_env = env_int("DLB200_PACK_CTAS_PER_SM", 0);
            if (u_env == 4)
                launch_tiles<FlatT, PACK, 4>(tab, flat, v_begin, v_end, scale, per_sm_env, stream);
            else
                launch_tiles<FlatT, PACK, 8>(tab, flat, v_begin, v_end, scale, per_sm_env, stream);
        } else {
            static bool attr_set[4] = {false, false, false, false};
            auto kern = pack_bulk_kernel<FlatT, PACK>;
            const int slot = (std::is_same<FlatT, float>::value ? 0 : 2) + (PACK ? 0 : 1);
            if (!attr_set[slot]) {
                DLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BulkSmem)));
                attr_set[slot] = true;
            }
            static const int bulk_per_sm = env_int("DLB200_BULK_CTAS_PER_SM", 3);
            const uint32_t tiles = (v_end - v_begin + kBulkTileBytes / 16 - 1) / (kBulkTileBytes / 16);
            const uint32_t grid = std::min<uint32_t>(tiles, (uint32_t)sm_count() * bulk_per_sm);
            kern<<<grid, kBulkThreads, sizeof(BulkSmem), stream>>>(tab, flat, v_begin, v_end, scale);
        }
    }
    DLB_CUDA(cudaGetLastError());
    return 0;
}

template <bool PACK>
int launch_move(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes, const int64_t* offsets,
                const int64_t* numels, int ntensors, float scale, int variant, cudaStream_t stream) {
    DLB_REQUIRE(ntensors >= 0, DLB_E_ARG, "dlb_%s: negative tensor count", PACK ? "pack" : "unpack");
    DLB_REQUIRE(flat_dtype == DLB_F32 || flat_dtype == DLB_BF16, DLB_E_ARG, "flat buffer dtype must be fp32 or bf16");
    DLB_REQUIRE(variant >= kAuto && variant <= kBulk, DLB_E_ARG, "pack variant %d (0 auto, 1 chunks, 2 tiles, 3 bulk)", variant);
    DLB_REQUIRE(ntensors == 0 || (tensors && tensor_dtypes && offsets && numels), DLB_E_ARG, "dlb_%s: null table", PACK ? "pack" : "unpack");
    // lists longer than one table are split into several launches
    for (int base = 0; base < ntensors; base += DLB_MAX_TENSORS) {
        const int n = ntensors - base < DLB_MAX_TENSORS ? ntensors - base : DLB_MAX_TENSORS;
        ChunkTable ct;
        bool aligned = false;
        uint32_t first = 0, end = 0;
        int rc = dlb_fill_table(ct.t, flat_dtype, tensors, tensor_dtypes, offsets, numels, base, n, &aligned, &first, &end);
        if (rc) return rc;
        uint32_t chunks = 0;
        bool same = true;
        for (int i = 0; i < ct.t.ntensors; ++i) {
            ct.chunk_prefix[i] = chunks;
            chunks += (ct.t.numel[i] + kChunk - 1) / kChunk;
            same = same && ct.t.dtype[i] == flat_dtype;
        }
        ct.chunk_prefix[ct.t.ntensors] = chunks;
        if (flat_dtype == DLB_F32)
            rc = launch_one<float, PACK>(ct, variant, aligned, same, first, end, (float*)flat, scale, stream);
        else
            rc = launch_one<__nv_bfloat16, PACK>(ct, variant, aligned, same, first, end, (__nv_bfloat16*)flat, scale, stream);
        if (rc) return rc;
    }
    return 0;
}

}  // namespace

// Fills `tab` with tensors [base, base+n) of the C-ABI arrays, sorted by offset.
int dlb_fill_table(TensorTable& tab, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                   const int64_t* offsets, const int64_t* numels, int base, int n, bool* vector_aligned,
                   uint32_t* first_elem, uint32_t* end_elem) {
    DLB_REQUIRE(n >= 0 && n <= DLB_MAX_TENSORS, DLB_E_LIMIT, "%d tensors in one table (at most %d)", n, DLB_MAX_TENSORS);
    int order[DLB_MAX_TENSORS];
    for (int i = 0; i < n; ++i) {
        const int j = base + i;
        DLB_REQUIRE(numels[j] >= 0 && numels[j] < (int64_t)UINT32_MAX && offsets[j] >= 0 && offsets[j] < (int64_t)UINT32_MAX &&
                        offsets[j] + numels[j] < (int64_t)UINT32_MAX,
                    DLB_E_LIMIT, "tensor %d: numel/offset beyond 2^32-1 elements", j);
        DLB_REQUIRE(tensor_dtypes[j] >= DLB_F32 && tensor_dtypes[j] <= DLB_F16, DLB_E_ARG, "tensor %d: bad dtype", j);
        order[i] = j;
    }
    std::stable_sort(order, order + n, [&](int a, int b) { return offsets[a] < offsets[b]; });
    const int per_vec = 16 / dlb_dtype_size(flat_dtype);
    bool aligned = true;
    int k = 0;
    uint32_t first = 0, end = 0;
    for (int i = 0; i < n; ++i) {
        const int j = order[i];
        if (numels[j] == 0) continue;  // nothing to move, and an empty entry would break the search
        const uint32_t off = (uint32_t)offsets[j], cnt = (uint32_t)numels[j];
        if (k == 0) first = off;
        aligned = aligned && off % per_vec == 0 && (k == 0 || off >= (end + per_vec - 1) / per_vec * per_vec);
        tab.tensor[k] = tensors[j];
        tab.offset[k] = off;
        tab.numel[k] = cnt;
        tab.dtype[k] = (uint8_t)tensor_dtypes[j];
        end = std::max(end, off + cnt);
        ++k;
    }
    tab.ntensors = k;
    tab.emu_stacked = 0;
    if (vector_aligned) *vector_aligned = aligned && k > 0;
    if (first_elem) *first_elem = first;
    if (end_elem) *end_elem = end;
    return 0;
}

extern "C" int dlb_pack(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                        const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream) {
    return launch_move<true>(flat, flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors, scale, kAuto,
                             (cudaStream_t)stream);
}

extern "C" int dlb_unpack(const void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                          const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream) {
    return launch_move<false>(const_cast<void*>(flat), flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors,
                              scale, kAuto, (cudaStream_t)stream);
}

extern "C" int dlb_pack_ex(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                           const int64_t* offsets, const int64_t* numels, int ntensors, float scale, int variant,
                           void* stream) {
    return launch_move<true>(flat, flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors, scale, variant,
                             (cudaStream_t)stream);
}

extern "C" int dlb_unpack_ex(const void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
                             const int64_t* offsets, const int64_t* numels, int ntensors, float scale, int variant,
                             void* stream) {
    return launch_move<false>(const_cast<void*>(flat), flat_dtype, tensors, tensor_dtypes, offsets, numels, ntensors,
                              scale, variant, (cudaStream_t)stream);
}
