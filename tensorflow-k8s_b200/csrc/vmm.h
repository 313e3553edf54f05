// vmm.h -- device memory that can be bound to an NVSwitch multicast object, and the plumbing that
// moves its handles between the ranks of a box (csrc/vmm.cu).  Used by the context when
// B200PS_AGG_MULTIMEM is selected; the default transport stays cudaMalloc + CUDA IPC.
#pragma once
#include <cuda.h>

#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

namespace b200ps {
namespace vmm {

bool available();  // the driver entry points resolve and the device reports multicast support

// one physical allocation of this device, mapped read/write at `va`
struct Block {
    CUmemGenericAllocationHandle handle = 0;
    size_t bytes = 0;
    CUdeviceptr va = 0;
    bool mapped = false;
};

size_t granularity(int device, size_t mc_bytes_hint, int n_devices);  // common multiple of the memory and multicast granularities
int alloc(int device, size_t bytes, Block* out);        // cuMemCreate (POSIX-fd exportable) + reserve + map + access
int export_fd(const Block& b, int* fd);
int import_fd(int device, int fd, size_t bytes, Block* out);  // a peer's allocation, mapped read/write here
void release(Block* b);

struct Multicast {
    CUmemGenericAllocationHandle handle = 0;
    size_t bytes = 0;
    CUdeviceptr va = 0;
    bool bound = false;
};
int mc_create(int n_devices, size_t bytes, Multicast* out, int* fd);  // rank 0
int mc_import(int fd, size_t bytes, Multicast* out);                   // other ranks
int mc_add_device(Multicast* mc, int device);                          // every rank, before anyone binds
int mc_bind_and_map(Multicast* mc, int device, const Block& local);    // every rank, after every rank has added its device
void mc_release(Multicast* mc);

// A tiny server that hands file descriptors to the other ranks over an abstract unix socket
// (SCM_RIGHTS): request byte 'A' -> the arena's fd, 'M' -> the multicast object's fd (rank 0).
class FdServer {
  public:
    ~FdServer() { stop(); }
    int start(const std::string& name, int arena_fd, int mc_fd);
    void stop();
    const std::string& name() const { return name_; }

  private:
    void run();
    std::string name_;
    int listen_fd_ = -1, arena_fd_ = -1, mc_fd_ = -1;
    std::atomic<bool> stop_{false};
    std::thread thread_;
};
int fetch_fd(const std::string& name, char what, int* fd);  // client side; retries while the peer comes up

}  // namespace vmm
}  // namespace b200ps
