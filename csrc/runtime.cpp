// runtime.cpp -- bring-up for the NVLink kernels: topology discovery, peer-mappable memory,
// handle exchange, multicast objects.  Host code only.
//
// The reference brings a job up by naming a NIC (GLOO_SOCKET_IFNAME, src/main.py:306-308) and
// moves tensors between processes by pickling them through mp.Queue (src/reducers.py:42-43).
// Here a job is one process per GPU on one NVSwitch box: each process allocates an arena,
// exports it (CUDA IPC handle, or a POSIX fd for VMM allocations that can also be bound to a
// multicast object), and maps every peer's arena.  The driver API is reached through
// cudaGetDriverEntryPoint, so this library has no link-time dependency on libcuda and can be
// loaded (and its host-only parts tested) on a machine without a GPU.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <errno.h>
#include <poll.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <string>

#include "../include/dlb200.h"

// ---- errors -------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void dlb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* dlb_last_error(void) { return g_err; }
extern "C" int dlb_version(void) { return 200; }

#define RT(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            dlb_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));  \
            return (int)e__;                                                                      \
        }                                                                                         \
    } while (0)

// ---- driver API through the runtime ---------------------------------------------------------------
template <typename Fn>
static int driver_fn(const char* name, Fn* out) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
        cudaGetLastError();
        dlb_set_error("driver entry point %s is not available (%s)", name, e != cudaSuccess ? cudaGetErrorString(e) : "not found");
        return DLB_E_UNSUPPORTED;
    }
    *out = reinterpret_cast<Fn>(p);
    return 0;
}

#define DRV_FN(var, name, ...)                      \
    CUresult (*var)(__VA_ARGS__) = nullptr;         \
    {                                               \
        int rc__ = driver_fn(name, &var);           \
        if (rc__) return rc__;                      \
    }

#define DRV(call)                                                                         \
    do {                                                                                  \
        CUresult r__ = (call);                                                            \
        if (r__ != CUDA_SUCCESS) {                                                        \
            dlb_set_error("%s:%d %s -> CUresult %d", __FILE__, __LINE__, #call, (int)r__); \
            return (int)r__;                                                              \
        }                                                                                 \
    } while (0)

static int ensure_context(int dev) {
    RT(cudaSetDevice(dev));
    RT(cudaFree(0));
    return 0;
}

// ---- plain allocations + legacy IPC -----------------------------------------------------------------
extern "C" int dlb_device_count(int* count) {
    RT(cudaGetDeviceCount(count));
    return 0;
}

extern "C" int dlb_can_access_peer(int dev, int peer, int* can) {
    RT(cudaDeviceCanAccessPeer(can, dev, peer));
    return 0;
}

extern "C" int dlb_enable_peer_access(int dev, int peer) {
    int cur = 0;
    RT(cudaGetDevice(&cur));
    RT(cudaSetDevice(dev));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    cudaSetDevice(cur);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    RT(e);
    return 0;
}

extern "C" int dlb_mem_alloc(size_t bytes, void** ptr) {
    RT(cudaMalloc(ptr, bytes));
    RT(cudaMemset(*ptr, 0, bytes));
    RT(cudaDeviceSynchronize());
    return 0;
}

extern "C" int dlb_mem_free(void* ptr) {
    RT(cudaFree(ptr));
    return 0;
}

extern "C" int dlb_ipc_export(void* ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    RT(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle, &h, 64);
    return 0;
}

extern "C" int dlb_ipc_open(const unsigned char handle[64], void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    RT(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int dlb_ipc_close(void* ptr) {
    RT(cudaIpcCloseMemHandle(ptr));
    return 0;
}

// ---- VMM allocations -------------------------------------------------------------------------------
static CUmemAllocationProp vmm_prop(int dev) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

extern "C" int dlb_vmm_granularity(int dev, int multicast, size_t* gran) {
    int rc = ensure_context(dev);
    if (rc) return rc;
    CUmemAllocationProp prop = vmm_prop(dev);
    DRV_FN(getGran, "cuMemGetAllocationGranularity", size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    size_t g = 0;
    DRV(getGran(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (multicast) {
        DRV_FN(mcGran, "cuMulticastGetGranularity", size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = 2;
        mp.size = g;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        DRV(mcGran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
        if (mg > g) g = mg;
    }
    *gran = g;
    return 0;
}

static int map_handle(int dev, CUmemGenericAllocationHandle h, size_t bytes, void** ptr) {
    DRV_FN(reserve, "cuMemAddressReserve", CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    DRV_FN(map, "cuMemMap", CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    DRV_FN(setAccess, "cuMemSetAccess", CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUdeviceptr va = 0;
    DRV(reserve(&va, bytes, 0, 0, 0));
    DRV(map(va, bytes, 0, h, 0));
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    DRV(setAccess(va, bytes, &acc, 1));
    *ptr = reinterpret_cast<void*>(va);
    return 0;
}

extern "C" int dlb_vmm_alloc(int dev, size_t bytes, uint64_t* handle, void** ptr, size_t* rounded) {
    size_t gran = 0;
    int rc = dlb_vmm_granularity(dev, 1, &gran);
    if (rc) {  // multicast granularity unavailable: fall back to the plain one
        rc = dlb_vmm_granularity(dev, 0, &gran);
        if (rc) return rc;
    }
    {   // the multicast granularity can grow with the object size: ask again for this size
        CUresult (*mcGran)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
        if (driver_fn("cuMulticastGetGranularity", &mcGran) == 0) {
            CUmulticastObjectProp mp;
            memset(&mp, 0, sizeof(mp));
            mp.numDevices = DLB_MAX_RANKS;
            mp.size = (bytes + gran - 1) / gran * gran;
            mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            size_t mg = 0;
            if (mcGran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
        }
    }
    const size_t size = (bytes + gran - 1) / gran * gran;
    CUmemAllocationProp prop = vmm_prop(dev);
    DRV_FN(create, "cuMemCreate", CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUmemGenericAllocationHandle h = 0;
    DRV(create(&h, size, &prop, 0));
    rc = map_handle(dev, h, size, ptr);
    if (rc) return rc;
    RT(cudaMemset(*ptr, 0, size));
    RT(cudaDeviceSynchronize());
    *handle = (uint64_t)h;
    *rounded = size;
    return 0;
}

extern "C" int dlb_vmm_export_fd(uint64_t handle, int* fd) {
    DRV_FN(exp, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    int out = -1;
    DRV(exp(&out, (CUmemGenericAllocationHandle)handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *fd = out;
    return 0;
}

extern "C" int dlb_vmm_import_fd(int fd, uint64_t* handle) {
    DRV_FN(imp, "cuMemImportFromShareableHandle", CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUmemGenericAllocationHandle h = 0;
    DRV(imp(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    *handle = (uint64_t)h;
    return 0;
}

extern "C" int dlb_vmm_map(int dev, uint64_t handle, size_t bytes, void** ptr) {
    int rc = ensure_context(dev);
    if (rc) return rc;
    return map_handle(dev, (CUmemGenericAllocationHandle)handle, bytes, ptr);
}

extern "C" int dlb_vmm_unmap(void* ptr, size_t bytes) {
    DRV_FN(unmap, "cuMemUnmap", CUdeviceptr, size_t);
    DRV_FN(afree, "cuMemAddressFree", CUdeviceptr, size_t);
    DRV(unmap((CUdeviceptr)ptr, bytes));
    DRV(afree((CUdeviceptr)ptr, bytes));
    return 0;
}

extern "C" int dlb_vmm_release(uint64_t handle) {
    DRV_FN(rel, "cuMemRelease", CUmemGenericAllocationHandle);
    DRV(rel((CUmemGenericAllocationHandle)handle));
    return 0;
}

// ---- multicast (NVLS) --------------------------------------------------------------------------------
extern "C" int dlb_mc_supported(int dev, int* yes) {
    *yes = 0;
    int ndev = 0;
    RT(cudaGetDeviceCount(&ndev));  // also initialises the driver; no context is needed below
    DRV_FN(getAttr, "cuDeviceGetAttribute", int*, CUdevice_attribute, CUdevice);
    DRV_FN(devGet, "cuDeviceGet", CUdevice*, int);
    CUdevice d;
    DRV(devGet(&d, dev));
    int v = 0;
    DRV(getAttr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d));
    *yes = v;
    return 0;
}

extern "C" int dlb_mc_create(int ndev, size_t bytes, uint64_t* mc_handle, size_t* rounded) {
    DRV_FN(mcGran, "cuMulticastGetGranularity", size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    DRV_FN(mcCreate, "cuMulticastCreate", CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)ndev;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    // dlb_vmm_alloc already rounded the arena to the recommended granularity for its size; only
    // the minimum is a hard requirement here, so an arena made by dlb_vmm_alloc is never re-rounded
    size_t g = 0;
    DRV(mcGran(&g, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    mp.size = (bytes + g - 1) / g * g;
    CUmemGenericAllocationHandle h = 0;
    DRV(mcCreate(&h, &mp));
    *mc_handle = (uint64_t)h;
    *rounded = mp.size;
    return 0;
}

extern "C" int dlb_mc_add_device(uint64_t mc_handle, int dev) {
    DRV_FN(add, "cuMulticastAddDevice", CUmemGenericAllocationHandle, CUdevice);
    DRV_FN(devGet, "cuDeviceGet", CUdevice*, int);
    CUdevice d;
    DRV(devGet(&d, dev));
    DRV(add((CUmemGenericAllocationHandle)mc_handle, d));
    return 0;
}

extern "C" int dlb_mc_bind(uint64_t mc_handle, size_t mc_offset, uint64_t mem_handle, size_t bytes) {
    DRV_FN(bind, "cuMulticastBindMem", CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
    DRV(bind((CUmemGenericAllocationHandle)mc_handle, mc_offset, (CUmemGenericAllocationHandle)mem_handle, 0, bytes, 0));
    return 0;
}

extern "C" int dlb_mc_map(int dev, uint64_t mc_handle, size_t bytes, void** mc_ptr) {
    int rc = ensure_context(dev);
    if (rc) return rc;
    return map_handle(dev, (CUmemGenericAllocationHandle)mc_handle, bytes, mc_ptr);
}

// ---- passing file descriptors between processes ------------------------------------------------------
static socklen_t abstract_addr(const char* name, sockaddr_un* addr) {
    memset(addr, 0, sizeof(*addr));
    addr->sun_family = AF_UNIX;
    const size_t n = strnlen(name, sizeof(addr->sun_path) - 2);
    memcpy(addr->sun_path + 1, name, n);  // leading NUL: abstract namespace, nothing to unlink
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

static long long now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (long long)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

extern "C" int dlb_fd_listen(const char* name, int* listen_fd) {
    int s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (s < 0) {
        dlb_set_error("socket: %s", strerror(errno));
        return DLB_E_SYS;
    }
    sockaddr_un addr;
    socklen_t len = abstract_addr(name, &addr);
    if (bind(s, (sockaddr*)&addr, len) != 0 || listen(s, 64) != 0) {
        dlb_set_error("bind/listen on '%s': %s", name, strerror(errno));
        close(s);
        return DLB_E_SYS;
    }
    *listen_fd = s;
    return 0;
}

extern "C" int dlb_fd_accept(int listen_fd, int* fd, int* tag, int timeout_ms) {
    pollfd p = {listen_fd, POLLIN, 0};
    int pr = poll(&p, 1, timeout_ms);
    if (pr <= 0) {
        dlb_set_error("waiting for a peer's descriptor: %s", pr == 0 ? "timed out" : strerror(errno));
        return pr == 0 ? DLB_E_TIMEOUT : DLB_E_SYS;
    }
    int c = accept(listen_fd, nullptr, nullptr);
    if (c < 0) {
        dlb_set_error("accept: %s", strerror(errno));
        return DLB_E_SYS;
    }
    int payload = -1;
    iovec iov = {&payload, sizeof(payload)};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    ssize_t n = recvmsg(c, &msg, 0);
    close(c);
    cmsghdr* cm = n > 0 ? CMSG_FIRSTHDR(&msg) : nullptr;
    if (cm == nullptr || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS) {
        dlb_set_error("recvmsg carried no descriptor (%s)", n < 0 ? strerror(errno) : "no SCM_RIGHTS");
        return DLB_E_SYS;
    }
    memcpy(fd, CMSG_DATA(cm), sizeof(int));
    *tag = payload;
    return 0;
}

extern "C" int dlb_fd_send(const char* name, int fd, int tag, int timeout_ms) {
    sockaddr_un addr;
    socklen_t len = abstract_addr(name, &addr);
    const long long deadline = now_ms() + timeout_ms;
    int s = -1;
    for (;;) {
        s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (s < 0) {
            dlb_set_error("socket: %s", strerror(errno));
            return DLB_E_SYS;
        }
        if (connect(s, (sockaddr*)&addr, len) == 0) break;
        close(s);
        if (now_ms() > deadline) {
            dlb_set_error("connect to '%s': %s (peer not listening)", name, strerror(errno));
            return DLB_E_TIMEOUT;
        }
        usleep(2000);
    }
    iovec iov = {&tag, sizeof(tag)};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg;
    memset(&msg, 0, sizeof(msg));
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(cm), &fd, sizeof(int));
    ssize_t n = sendmsg(s, &msg, 0);
    const int saved = errno;
    close(s);
    if (n != (ssize_t)sizeof(tag)) {
        dlb_set_error("sendmsg to '%s': %s", name, strerror(saved));
        return DLB_E_SYS;
    }
    return 0;
}

extern "C" int dlb_fd_close(int fd) { return close(fd) == 0 ? 0 : DLB_E_SYS; }

// ---- topology -----------------------------------------------------------------------------------------
namespace {
struct Nvml {
    void* lib = nullptr;
    int (*init)() = nullptr;
    int (*byPci)(const char*, void**) = nullptr;
    int (*linkState)(void*, unsigned, unsigned*) = nullptr;
    int (*linkVersion)(void*, unsigned, unsigned*) = nullptr;
    bool ok = false;
    Nvml() {
        lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_LOCAL);
        if (!lib) return;
        init = (int (*)())dlsym(lib, "nvmlInit_v2");
        byPci = (int (*)(const char*, void**))dlsym(lib, "nvmlDeviceGetHandleByPciBusId_v2");
        linkState = (int (*)(void*, unsigned, unsigned*))dlsym(lib, "nvmlDeviceGetNvLinkState");
        linkVersion = (int (*)(void*, unsigned, unsigned*))dlsym(lib, "nvmlDeviceGetNvLinkVersion");
        ok = init && byPci && linkState && init() == 0;
    }
};

void append(std::string& s, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    s += buf;
}
}  // namespace

extern "C" int dlb_topology_json(char* out, size_t cap) {
    if (!out || cap == 0) return DLB_E_ARG;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    int rtv = 0, drv = 0;
    cudaRuntimeGetVersion(&rtv);
    cudaDriverGetVersion(&drv);
    static Nvml nvml;
    std::string s;
    append(s, "{\"runtime\": %d, \"driver\": %d, \"nvml\": %s, \"devices\": [", rtv, drv, nvml.ok ? "true" : "false");
    for (int d = 0; d < n; ++d) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, d) != cudaSuccess) continue;
        char bus[32] = "";
        cudaDeviceGetPCIBusId(bus, sizeof(bus), d);
        int mc = 0;
        dlb_mc_supported(d, &mc);
        int links = -1, version = -1;
        if (nvml.ok) {
            void* h = nullptr;
            if (nvml.byPci(bus, &h) == 0) {
                links = 0;
                for (unsigned l = 0; l < 18; ++l) {
                    unsigned active = 0;
                    if (nvml.linkState(h, l, &active) == 0 && active) {
                        ++links;
                        unsigned v = 0;
                        if (version < 0 && nvml.linkVersion && nvml.linkVersion(h, l, &v) == 0) version = (int)v;
                    }
                }
            }
        }
        append(s, "%s{\"index\": %d, \"name\": \"%s\", \"cc\": \"%d.%d\", \"sms\": %d, \"mem_gib\": %.1f, \"pci\": \"%s\", "
                  "\"l2_mib\": %.0f, \"multicast\": %d, \"nvlinks_active\": %d, \"nvlink_version\": %d}",
               d ? ", " : "", d, p.name, p.major, p.minor, p.multiProcessorCount,
               (double)p.totalGlobalMem / (1024.0 * 1024.0 * 1024.0), bus, (double)p.l2CacheSize / (1024.0 * 1024.0), mc,
               links, version);
    }
    s += "], \"peer_access\": [";
    for (int a = 0; a < n; ++a) {
        s += a ? ", [" : "[";
        for (int b = 0; b < n; ++b) {
            int can = a == b;
            if (a != b) cudaDeviceCanAccessPeer(&can, a, b);
            append(s, "%s%d", b ? ", " : "", can);
        }
        s += "]";
    }
    s += "], \"native_atomics\": [";
    for (int a = 0; a < n; ++a) {
        s += a ? ", [" : "[";
        for (int b = 0; b < n; ++b) {
            int v = a == b;
            if (a != b) cudaDeviceGetP2PAttribute(&v, cudaDevP2PAttrNativeAtomicSupported, a, b);
            append(s, "%s%d", b ? ", " : "", v);
        }
        s += "]";
    }
    s += "], \"performance_rank\": [";
    for (int a = 0; a < n; ++a) {
        s += a ? ", [" : "[";
        for (int b = 0; b < n; ++b) {
            int v = 0;
            if (a != b) cudaDeviceGetP2PAttribute(&v, cudaDevP2PAttrPerformanceRank, a, b);
            append(s, "%s%d", b ? ", " : "", v);
        }
        s += "]";
    }
    s += "]}";
    cudaGetLastError();
    snprintf(out, cap, "%s", s.c_str());
    return 0;
}
