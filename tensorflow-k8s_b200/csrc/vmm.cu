// vmm.cu -- virtual-memory-managed device allocations, NVSwitch multicast objects and the exchange of
// their POSIX file-descriptor handles between the ranks of one box (vmm.h).
//
// Reference: none.  This is the transport under B200PS_AGG_MULTIMEM: every rank's arena is a
// cuMemCreate allocation (exportable as a file descriptor), the first 2 * P_pad floats of every arena
// -- replica and partial sum -- are bound to one multicast object, and a kernel that addresses the
// multicast mapping reads the SUM of all ranks' partial sums with one multimem.ld_reduce (added inside
// the switch) and writes every replica with one multimem.st.  Driver entry points are resolved through
// the runtime (cudaGetDriverEntryPoint), so the library does not link libcuda.
#include "vmm.h"

#include <cuda_runtime.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <chrono>
#include <cstring>

#include "common.h"

namespace b200ps {
namespace vmm {
namespace {

struct Driver {
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
    bool ok = false;
};

template <typename F>
bool resolve(const char* name, F* out) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || !p) return false;
    *out = reinterpret_cast<F>(p);
    return true;
}

const Driver& drv() {
    static Driver d = [] {
        Driver x;
        x.ok = resolve("cuMemCreate", &x.MemCreate) && resolve("cuMemRelease", &x.MemRelease) &&
               resolve("cuMemAddressReserve", &x.MemAddressReserve) && resolve("cuMemAddressFree", &x.MemAddressFree) &&
               resolve("cuMemMap", &x.MemMap) && resolve("cuMemUnmap", &x.MemUnmap) && resolve("cuMemSetAccess", &x.MemSetAccess) &&
               resolve("cuMemExportToShareableHandle", &x.MemExportToShareableHandle) &&
               resolve("cuMemImportFromShareableHandle", &x.MemImportFromShareableHandle) &&
               resolve("cuMemGetAllocationGranularity", &x.MemGetAllocationGranularity) &&
               resolve("cuMulticastCreate", &x.MulticastCreate) && resolve("cuMulticastAddDevice", &x.MulticastAddDevice) &&
               resolve("cuMulticastBindMem", &x.MulticastBindMem) && resolve("cuMulticastGetGranularity", &x.MulticastGetGranularity) &&
               resolve("cuDeviceGetAttribute", &x.DeviceGetAttribute) && resolve("cuGetErrorString", &x.GetErrorString);
        return x;
    }();
    return d;
}

const char* err_str(CUresult r) {
    const char* s = nullptr;
    if (drv().GetErrorString && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown driver error";
}

#define B200PS_DRV(expr)                                                                                    \
    do {                                                                                                    \
        CUresult _r = (expr);                                                                               \
        if (_r != CUDA_SUCCESS)                                                                             \
            return ::b200ps::fail(B200PS_ERR_CUDA, "%s: %s (%s:%d)", #expr, err_str(_r), __FILE__, __LINE__); \
    } while (0)

CUmemAllocationProp mem_prop(int device) {
    CUmemAllocationProp p{};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = device;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

int map_rw(int device, CUmemGenericAllocationHandle h, size_t bytes, CUdeviceptr* va) {
    const Driver& d = drv();
    B200PS_DRV(d.MemAddressReserve(va, bytes, 0, 0, 0));
    B200PS_DRV(d.MemMap(*va, bytes, 0, h, 0));
    CUmemAccessDesc acc{};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    B200PS_DRV(d.MemSetAccess(*va, bytes, &acc, 1));
    return B200PS_OK;
}

}  // namespace

bool available() {
    const Driver& d = drv();
    if (!d.ok) return false;
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return false;
    return v != 0;
}

size_t granularity(int device, size_t mc_bytes_hint, int n_devices) {
    const Driver& d = drv();
    size_t g = 2u << 20, mg = 0;
    const CUmemAllocationProp p = mem_prop(device);
    if (d.ok && d.MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) g = 2u << 20;
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)n_devices;
    mp.size = mc_bytes_hint;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (d.ok && d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
    return g;
}

int alloc(int device, size_t bytes, Block* out) {
    const Driver& d = drv();
    B200PS_REQUIRE(d.ok, B200PS_ERR_CUDA, "the CUDA driver does not offer the virtual memory management entry points");
    const CUmemAllocationProp p = mem_prop(device);
    B200PS_DRV(d.MemCreate(&out->handle, bytes, &p, 0));
    out->bytes = bytes;
    if (int rc = map_rw(device, out->handle, bytes, &out->va)) return rc;
    out->mapped = true;
    return B200PS_OK;
}

int export_fd(const Block& b, int* fd) {
    B200PS_DRV(drv().MemExportToShareableHandle(fd, b.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    return B200PS_OK;
}

int import_fd(int device, int fd, size_t bytes, Block* out) {
    const Driver& d = drv();
    B200PS_DRV(d.MemImportFromShareableHandle(&out->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    out->bytes = bytes;
    if (int rc = map_rw(device, out->handle, bytes, &out->va)) return rc;
    out->mapped = true;
    return B200PS_OK;
}

void release(Block* b) {
    const Driver& d = drv();
    if (!d.ok) return;
    if (b->mapped) {
        d.MemUnmap(b->va, b->bytes);
        d.MemAddressFree(b->va, b->bytes);
    }
    if (b->handle) d.MemRelease(b->handle);
    *b = Block{};
}

int mc_create(int n_devices, size_t bytes, Multicast* out, int* fd) {
    const Driver& d = drv();
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)n_devices;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    B200PS_DRV(d.MulticastCreate(&out->handle, &mp));
    out->bytes = bytes;
    B200PS_DRV(d.MemExportToShareableHandle(fd, out->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    return B200PS_OK;
}

int mc_import(int fd, size_t bytes, Multicast* out) {
    B200PS_DRV(drv().MemImportFromShareableHandle(&out->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    out->bytes = bytes;
    return B200PS_OK;
}

int mc_add_device(Multicast* mc, int device) {
    B200PS_DRV(drv().MulticastAddDevice(mc->handle, device));
    return B200PS_OK;
}

int mc_bind_and_map(Multicast* mc, int device, const Block& local) {
    const Driver& d = drv();
    B200PS_REQUIRE(local.bytes >= mc->bytes, B200PS_ERR_RANGE, "the arena is smaller than the multicast window");
    B200PS_DRV(d.MulticastBindMem(mc->handle, 0, local.handle, 0, mc->bytes, 0));
    if (int rc = map_rw(device, mc->handle, mc->bytes, &mc->va)) return rc;
    mc->bound = true;
    return B200PS_OK;
}

void mc_release(Multicast* mc) {
    const Driver& d = drv();
    if (!d.ok) return;
    if (mc->va) {
        d.MemUnmap(mc->va, mc->bytes);
        d.MemAddressFree(mc->va, mc->bytes);
    }
    if (mc->handle) d.MemRelease(mc->handle);
    *mc = Multicast{};
}

// ---- file descriptors over an abstract unix socket -------------------------------------------------
namespace {

socklen_t make_addr(const std::string& name, sockaddr_un* a) {
    memset(a, 0, sizeof *a);
    a->sun_family = AF_UNIX;
    const size_t n = name.size() < sizeof(a->sun_path) - 2 ? name.size() : sizeof(a->sun_path) - 2;
    memcpy(a->sun_path + 1, name.data(), n);  // sun_path[0] = 0: abstract namespace, nothing to unlink
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

bool send_fd(int sock, int fd) {
    char byte = fd >= 0 ? 'k' : 'n';
    iovec io{&byte, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof ctrl);
    msghdr msg{};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    if (fd >= 0) {
        msg.msg_control = ctrl;
        msg.msg_controllen = sizeof ctrl;
        cmsghdr* c = CMSG_FIRSTHDR(&msg);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int));
        memcpy(CMSG_DATA(c), &fd, sizeof(int));
    }
    return sendmsg(sock, &msg, 0) == 1;
}

int recv_fd(int sock) {
    char byte = 0;
    iovec io{&byte, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr msg{};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    if (recvmsg(sock, &msg, 0) != 1 || byte != 'k') return -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c))
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
            int fd = -1;
            memcpy(&fd, CMSG_DATA(c), sizeof(int));
            return fd;
        }
    return -1;
}

}  // namespace

int FdServer::start(const std::string& name, int arena_fd, int mc_fd) {
    name_ = name;
    arena_fd_ = arena_fd;
    mc_fd_ = mc_fd;
    listen_fd_ = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    B200PS_REQUIRE(listen_fd_ >= 0, B200PS_ERR_INVALID, "socket(): %s", strerror(errno));
    sockaddr_un a;
    const socklen_t len = make_addr(name, &a);
    if (bind(listen_fd_, (sockaddr*)&a, len) != 0 || listen(listen_fd_, 32) != 0) {
        const int e = errno;
        close(listen_fd_);
        listen_fd_ = -1;
        return fail(B200PS_ERR_INVALID, "cannot serve handles on the abstract socket '%s': %s", name.c_str(), strerror(e));
    }
    stop_ = false;
    thread_ = std::thread([this] { run(); });
    return B200PS_OK;
}

void FdServer::run() {
    while (!stop_) {
        pollfd p{listen_fd_, POLLIN, 0};
        if (poll(&p, 1, 100) <= 0) continue;
        const int c = accept4(listen_fd_, nullptr, nullptr, SOCK_CLOEXEC);
        if (c < 0) continue;
        char what = 0;
        pollfd q{c, POLLIN, 0};
        if (poll(&q, 1, 2000) > 0 && recv(c, &what, 1, 0) == 1) send_fd(c, what == 'A' ? arena_fd_ : what == 'M' ? mc_fd_ : -1);
        close(c);
    }
}

void FdServer::stop() {
    stop_ = true;
    if (thread_.joinable()) thread_.join();
    if (listen_fd_ >= 0) close(listen_fd_);
    listen_fd_ = -1;
    if (arena_fd_ >= 0) close(arena_fd_);
    if (mc_fd_ >= 0) close(mc_fd_);
    arena_fd_ = mc_fd_ = -1;
}

int fetch_fd(const std::string& name, char what, int* fd) {
    sockaddr_un a;
    const socklen_t len = make_addr(name, &a);
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        const int s = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        B200PS_REQUIRE(s >= 0, B200PS_ERR_INVALID, "socket(): %s", strerror(errno));
        if (connect(s, (sockaddr*)&a, len) == 0) {
            int got = -1;
            if (send(s, &what, 1, 0) == 1) got = recv_fd(s);
            close(s);
            B200PS_REQUIRE(got >= 0, B200PS_ERR_INVALID, "the rank serving '%s' did not hand over handle '%c'", name.c_str(), what);
            *fd = got;
            return B200PS_OK;
        }
        close(s);
        if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(30))
            return fail(B200PS_ERR_INVALID, "no rank is serving handles on '%s' (waited 30 s): %s", name.c_str(), strerror(errno));
        usleep(20000);
    }
}

}  // namespace vmm
}  // namespace b200ps
