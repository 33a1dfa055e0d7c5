/*
 * dlb200.h -- C ABI of libdlb200.so, the native half of distributed_learning_b200.
 *
 * Everything on the gradient path of the framework that runs on the GPU is behind these entry
 * points: multi-tensor pack / unpack (gather + scale + cast into a fusion buffer and back),
 * the fused SGD-momentum update, the NVLink allreduce kernels (one-shot, two-shot, NVLS
 * multicast) with their device-side barrier, and the bring-up that the kernels need (topology
 * discovery, peer-mappable allocations, handle export/import, multicast objects).
 *
 * Conventions
 *   - Every function returns 0 on success.  A positive value is a cudaError_t / CUresult, a
 *     negative value one of DLB_E_*; dlb_last_error() returns a message for the calling thread.
 *   - Pointers named *_peers / bufs / sigs are host arrays of `nranks` device pointers, index =
 *     rank; entry [rank] is the caller's own memory, the others are peer mappings.
 *   - `stream` is a cudaStream_t passed as void*.  Nothing here synchronises the host unless
 *     its comment says so; nothing allocates device memory except dlb_mem_* / dlb_mc_*.
 *   - dtype codes: DLB_F32, DLB_BF16, DLB_F16.
 *
 * The reference (zdule/distributed_learning) has no native code; the citations below name the
 * Python it replaces.
 */
#ifndef DLB200_H
#define DLB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLB_MAX_RANKS 8     /* one NVSwitch domain */
#define DLB_MAX_BLOCKS 64   /* CTAs of one communication kernel */
#define DLB_MAX_TENSORS 512 /* tensors per multi-tensor launch (11 KB of kernel parameters; longer lists become several launches) */
#define DLB_SIGNAL_BYTES (2 * DLB_MAX_BLOCKS * DLB_MAX_RANKS * 4) /* size of one rank's signal pad */
#define DLB_STATE_WORDS (DLB_MAX_BLOCKS + 8)                       /* local epoch counters + error word */

enum { DLB_F32 = 0, DLB_BF16 = 1, DLB_F16 = 2 };

enum {
    DLB_E_ARG = -1,         /* bad argument */
    DLB_E_UNSUPPORTED = -2, /* not available on this system (e.g. no multicast) */
    DLB_E_SYS = -3,         /* a system call failed (socket, fd passing) */
    DLB_E_TIMEOUT = -4,     /* a device-side wait gave up: a peer never arrived */
    DLB_E_LIMIT = -5        /* a size is beyond a documented maximum */
};

const char* dlb_last_error(void);
int dlb_version(void);

/* ---- topology discovery (replaces the NIC name `i` of reference src/main.py:306-308) ---------
 * Writes a JSON document into `out` (NUL-terminated, truncated to `cap`): device names, SM
 * counts, memory, the peer-access / native-atomics / performance-rank matrices, NVLink link
 * states from NVML when libnvidia-ml is loadable, and multicast (NVLS) support. */
int dlb_topology_json(char* out, size_t cap);
int dlb_device_count(int* count);
int dlb_can_access_peer(int dev, int peer, int* can);
int dlb_enable_peer_access(int dev, int peer); /* single-process multi-GPU mode */

/* ---- peer-mappable memory -----------------------------------------------------------------
 * dlb_mem_alloc: cudaMalloc'd, zero-filled block on the current device.
 * dlb_ipc_export/open/close: legacy CUDA IPC, 64-byte handles, for one-process-per-GPU.
 * The replaced mechanism is torch.multiprocessing's tensor pickling through mp.Queue
 * (reference src/reducers.py:42-43). */
int dlb_mem_alloc(size_t bytes, void** ptr);
int dlb_mem_free(void* ptr);
int dlb_ipc_export(void* ptr, unsigned char handle[64]);
int dlb_ipc_open(const unsigned char handle[64], void** ptr);
int dlb_ipc_close(void* ptr);

/* VMM allocations that can be shared as POSIX file descriptors and bound to a multicast
 * object (NVLS).  dlb_vmm_alloc rounds `bytes` up to the granularity and reports it.
 * dlb_fd_listen / dlb_fd_accept / dlb_fd_send pass a descriptor (plus an int tag, e.g. the
 * sender's rank) between processes over an abstract-namespace unix socket named `name`
 * (SCM_RIGHTS): every rank listens, a barrier follows, every rank sends to every peer, then
 * accepts nranks-1 descriptors. */
int dlb_vmm_granularity(int dev, int multicast, size_t* gran);
int dlb_vmm_alloc(int dev, size_t bytes, uint64_t* handle, void** ptr, size_t* rounded);
int dlb_vmm_export_fd(uint64_t handle, int* fd);
int dlb_vmm_import_fd(int fd, uint64_t* handle);
int dlb_vmm_map(int dev, uint64_t handle, size_t bytes, void** ptr);
int dlb_vmm_unmap(void* ptr, size_t bytes);
int dlb_vmm_release(uint64_t handle);
int dlb_fd_listen(const char* name, int* listen_fd);
int dlb_fd_accept(int listen_fd, int* fd, int* tag, int timeout_ms);
int dlb_fd_send(const char* name, int fd, int tag, int timeout_ms);
int dlb_fd_close(int fd);

/* Multicast objects.  Rank 0 creates and exports, everyone imports, adds its device, binds its
 * VMM allocation at `mc_offset`, and maps the multicast address range. */
int dlb_mc_supported(int dev, int* yes);
int dlb_mc_create(int ndev, size_t bytes, uint64_t* mc_handle, size_t* rounded);
int dlb_mc_add_device(uint64_t mc_handle, int dev);
int dlb_mc_bind(uint64_t mc_handle, size_t mc_offset, uint64_t mem_handle, size_t bytes);
int dlb_mc_map(int dev, uint64_t mc_handle, size_t bytes, void** mc_ptr);

/* ---- host-side synthetic input (reference random_data_generator, src/data.py:129-132) -------
 * Fill `numel` elements at `dst` (host memory, normally the pinned slot an H2D copy will read)
 * with uniform values in [0,1) (fp32 or bf16) or labels in [0,high).  Counter based: element i of
 * batch `batch_index` depends only on (seed, batch_index, i), so the result is independent of
 * `nthreads`, which only sets how many threads share the work (the call returns when all are
 * done).  The reference draws the same data from one serial host RNG stream, 69 ms per
 * 128x3x224x224 batch on the B200 host. */
int dlb_host_fill_uniform(void* dst, int dtype, int64_t numel, uint64_t seed, uint64_t batch_index, int nthreads);
int dlb_host_fill_randint(int64_t* dst, int64_t numel, int64_t high, uint64_t seed, uint64_t batch_index);

/* ---- K-pack / K-unpack (reference Group.fuse / Group.unfuse, src/ourdist.py:28-43) -----------
 * One launch moves up to DLB_MAX_TENSORS tensors (longer lists: several launches) between their own storage and a flat fusion
 * buffer: flat[offset[i] + k] = cast(tensor_i[k] * scale)  (pack)  or the reverse (unpack).
 * `tensor_dtypes[i]` is the dtype of tensor i, `flat_dtype` that of the buffer; offsets and
 * counts are in elements.  A NULL tensor pointer packs zeros (a parameter without gradient). */
int dlb_pack(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
             const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream);
int dlb_unpack(const void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes,
               const int64_t* offsets, const int64_t* numels, int ntensors, float scale, void* stream);
/* The same with the kernel named: 0 = pick (what dlb_pack does), 1 = "chunks" (one CTA per 8192
 * elements of one tensor; any layout), 2 = "tiles" (persistent CTAs over equal tiles of the flat
 * buffer, owner of each 16-byte vector looked up in shared memory), 3 = "bulk" (cp.async.bulk +
 * mbarrier through a shared-memory ring; tensor dtype must equal the flat dtype).  2 and 3 need a
 * vector-aligned layout: ascending non-overlapping offsets, each a multiple of 16 bytes of the flat
 * dtype, and a flat buffer that extends to the next 16-byte boundary after its last tensor (the pad
 * elements are written as zeros by pack). */
int dlb_pack_ex(void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes, const int64_t* offsets,
                const int64_t* numels, int ntensors, float scale, int variant, void* stream);
int dlb_unpack_ex(const void* flat, int flat_dtype, void* const* tensors, const int* tensor_dtypes, const int64_t* offsets,
                  const int64_t* numels, int ntensors, float scale, int variant, void* stream);

/* ---- K-sgd (reference optim.SGD(lr=0.01, momentum=0.5), src/main.py:36,91) -------------------
 * buf = momentum*buf + (1-dampening)*(g*grad_scale + weight_decay*p);  p -= lr*(nesterov ? g'+momentum*buf : buf)
 * on flat arrays: fp32 master parameters and momentum, gradient in `grad_dtype`, and an
 * optional bf16 copy of the new parameters (`shadow`, may be NULL) for bf16 models. */
typedef struct {
    float lr, momentum, dampening, weight_decay, grad_scale;
    int nesterov;
} dlb_sgd_hyper_t;

int dlb_sgd_flat(float* param, float* momentum_buf, const void* grad, int grad_dtype, void* shadow_bf16,
                 int64_t numel, const dlb_sgd_hyper_t* hyper, void* stream);
/* multi-tensor form for parameters that are not flattened; a NULL gradient skips the tensor */
int dlb_sgd_multi(void* const* params, void* const* momentum_bufs, void* const* grads, const int* grad_dtypes,
                  const int64_t* numels, int ntensors, const dlb_sgd_hyper_t* hyper, void* stream);

/* ---- K-ar: in-kernel allreduce over peer memory (replaces reference src/allreduce.py) --------
 * Common arguments
 *   bufs    [nranks] the same fusion buffer in every rank's arena (peer mapped), `numel`
 *           elements of `dtype`, 16-byte aligned, numel a multiple of 16 bytes' worth
 *   sigs    [nranks] signal pads of DLB_SIGNAL_BYTES each (peer mapped, zero at start)
 *   state   local, DLB_STATE_WORDS uint32, zero at start: per-CTA epoch counters + error word
 *   scale   applied to the sum (1/nranks for an average, or 1 if the pack pre-scaled)
 *   nblocks CTAs to use (<= DLB_MAX_BLOCKS); every rank must pass the same value, and every
 *           rank must issue the same sequence of K-ar launches
 *   sgd     optional fused epilogue: when non-NULL the reduced gradient is consumed in place by
 *           the SGD update of `sgd_param/sgd_mom/sgd_shadow` (flat, same layout as the buffer)
 *
 * dlb_allreduce_twoshot: reduce-scatter by pulling slice `rank` from every peer, summing in fp32
 *   in rank order, and pushing the result into every peer's buffer; in place; two barriers.
 * dlb_allreduce_oneshot: every rank pulls everything from every peer and writes the sum to its
 *   local `out` (may not alias bufs[rank]); one barrier (+ a trailing one if end_barrier).
 * dlb_allreduce_nvls: multimem.ld_reduce of slice `rank` through the switch and multimem.st of
 *   the result; `mc_buf` is this rank's multicast mapping of the same buffer.
 */
typedef struct {
    float* param;        /* fp32 master parameters, flat */
    float* momentum_buf; /* fp32 momentum, flat */
    void* shadow_bf16;   /* optional bf16 copy of the parameters */
    dlb_sgd_hyper_t hyper;
} dlb_fused_sgd_t;

int dlb_allreduce_twoshot(void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks,
                          int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                          const dlb_fused_sgd_t* sgd, void* stream);
int dlb_allreduce_oneshot(void* const* bufs, void* out, void* const* sigs, uint32_t* state, int rank, int nranks,
                          int64_t numel, int dtype, float scale, int nblocks, int nthreads, int end_barrier,
                          const dlb_fused_sgd_t* sgd, void* stream);
int dlb_allreduce_nvls(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks,
                       int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                       const dlb_fused_sgd_t* sgd, void* stream);
/* dlb_allreduce_nvls with eight multimem.ld_reduce in flight per thread instead of four (same
 * arguments, same result; dlb_allreduce's `unroll`). */
int dlb_allreduce_nvls_deep(void* mc_buf, void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks,
                            int64_t numel, int dtype, float scale, int nblocks, int nthreads,
                            const dlb_fused_sgd_t* sgd, void* stream);
/* The general form: everything above plus the FUSED PACK PROLOGUE, the unroll of the nvls variant
 * and the phase trace, in one descriptor (zero-initialise it, then set what you need).
 *   tensors/tensor_dtypes/offsets/numels/ntensors  when ntensors > 0 the kernel first gathers the
 *           bucket's gradient tensors into bufs[rank] (flat[offset_i + k] = cast(tensor_i[k] *
 *           pack_scale); a NULL tensor packs zeros), each CTA packing exactly the tiles its peers'
 *           same-numbered CTAs will pull, then runs the entry barrier: one launch per bucket
 *           instead of dlb_pack + allreduce (reference Group.fuse, src/ourdist.py:28-36, followed
 *           by the reducer).  Needs the vector-aligned layout of dlb_pack_ex and at most
 *           DLB_MAX_TENSORS tensors.
 *   unroll  nvls only: switch reductions in flight per thread, 2 / 4 / 8 (0 = 4).  Every rank must
 *           pass the same value (it sets the tile that CTAs own).
 *   trace   optional device pointer to 8 uint64: %globaltimer (ns) of CTA 0 at start, packed, entry
 *           barrier passed, data phase done, exit barrier passed, epilogue done. */
enum {
    DLB_ALGO_TWOSHOT = 0,
    DLB_ALGO_ONESHOT = 1,
    DLB_ALGO_NVLS = 2,
    DLB_ALGO_TWOSHOT_BULK = 3 /* two-shot with the bytes moved by the bulk-copy engine (cp.async.bulk + mbarrier through a
                                 shared-memory ring; 512 threads per CTA); same result as DLB_ALGO_TWOSHOT */
};

typedef struct {
    int algo, rank, nranks;
    void* const* bufs;
    void* out;    /* one-shot */
    void* mc_buf; /* nvls */
    void* const* sigs;
    uint32_t* state;
    int64_t numel;
    int dtype;
    float scale;
    int nblocks, nthreads, unroll, end_barrier;
    const dlb_fused_sgd_t* sgd;
    void* const* tensors;
    const int* tensor_dtypes;
    const int64_t* offsets;
    const int64_t* numels;
    int ntensors;
    float pack_scale; /* 0 is read as 1 */
    unsigned long long* trace;
} dlb_allreduce_desc_t;

int dlb_allreduce(const dlb_allreduce_desc_t* desc, void* stream);

/* Every rank copies `nbytes` (multiple of 16) of bufs[root] into its own bufs[rank]; two barriers, so
 * the root may refill its buffer as soon as the kernel is done.  Gives every replica rank 0's initial
 * parameters without another communication library (the reference relies on equal seeds,
 * src/main.py:31-34). */
int dlb_broadcast(void* const* bufs, void* const* sigs, uint32_t* state, int rank, int nranks, int root, int64_t nbytes,
                  int nblocks, int nthreads, void* stream);
/* device-side barrier on its own (used at bring-up and by tests) */
int dlb_barrier(void* const* sigs, uint32_t* state, int rank, int nranks, int nblocks, void* stream);
/* spin budget of the device-side waits in milliseconds (default 600000, or $DLB200_WAIT_TIMEOUT_MS; 0 = wait forever) */
int dlb_set_wait_timeout_ms(int ms);
/* reads the error word of `state` (synchronises `stream`); 0 or DLB_E_TIMEOUT */
int dlb_comm_check(uint32_t* state, void* stream);

/* Single-GPU emulation of `nranks` ranks for tests: ONE cooperative launch in which
 * blockIdx.y is the rank, so the ranks' CTAs are guaranteed co-resident (separate launches that
 * wait on each other on one GPU are not).  All buffers live on the current device.
 * algo: 0 = two-shot, 1 = one-shot (outs[r] receives rank r's result).  `sgds` (NULL or nranks
 * entries) enables the fused SGD epilogue per emulated rank.
 * `flags` is scratch: nranks * (DLB_SIGNAL_BYTES + 4*DLB_STATE_WORDS) bytes, zero before first use. */
int dlb_allreduce_emulated(int algo, void* const* bufs, void* const* outs, const dlb_fused_sgd_t* sgds, void* flags,
                           int nranks, int64_t numel, int dtype, float scale, int nblocks, int nthreads, void* stream);
size_t dlb_emulated_flags_bytes(int nranks);
/* The emulation in descriptor form, with the fused pack prologue: tensor i of emulated rank r is
 * read at tensors[i] + r * numels[i] * sizeof(dtype_i), i.e. each tensor is stored as [nranks, numel]. */
typedef struct {
    int algo, nranks;
    void* const* bufs;
    void* const* outs;
    const dlb_fused_sgd_t* sgds;
    void* flags;
    int64_t numel;
    int dtype;
    float scale;
    int nblocks, nthreads;
    void* const* tensors;
    const int* tensor_dtypes;
    const int64_t* offsets;
    const int64_t* numels;
    int ntensors;
    float pack_scale;
    unsigned long long* trace; /* 8 uint64 per emulated rank */
} dlb_emulated_desc_t;
int dlb_allreduce_emulated_ex(const dlb_emulated_desc_t* desc, void* stream);

/* dlb_allreduce_emulated without the caller-provided scratch, for callers that cannot allocate
 * device memory (a plug-in inside a process whose allocator it does not own, a self-test on a
 * full device, a bare C host).  No counterpart in the reference: it has no native code and no
 * notion of scratch memory.  The flags live in a fixed array inside the library
 * (dlb_emulated_static_scratch_bytes() = 35 072 bytes, resident on every GPU that loads the
 * library until the process exits), sized for DLB_MAX_RANKS ranks and DLB_MAX_BLOCKS CTAs per
 * rank; larger requests return DLB_E_LIMIT and launch nothing.  The first call on a device looks
 * up the array's device address, so make it before any stream capture on that device.  Every call
 * clears the part of the array it uses on `stream` before the kernel.  All calls on a device share
 * the one array: calls whose GPU work overlaps race, so serialise them.  Inputs and outputs are
 * the caller's as before; nothing is allocated or freed.  No fused-SGD option. */
int dlb_allreduce_emulated_static(int algo, void* const* bufs, void* const* outs, int nranks, int64_t numel,
                                  int dtype, float scale, int nblocks, int nthreads, void* stream);
size_t dlb_emulated_static_scratch_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* DLB200_H */
