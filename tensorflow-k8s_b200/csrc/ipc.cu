// ipc.cu -- device buffers that other ranks of the box can address (include/b200ps.h
// "Transport group").
//
// Reference: none -- between PS and WORKER tasks (k8s-operator.md:6) a TFJob moves tensors with
// TensorFlow's gRPC runtime over the pod network.  On one NVSwitch box the transport is the
// memory system itself: every rank (one process per GPU) exports its gradient and parameter
// buffers as CUDA IPC handles, every other rank maps them, and the step kernel (dense.cu) loads
// gradients from and stores parameters to the mapped peer pointers over NVLink.
#include <cuda_runtime.h>

#include <cstring>

#include "common.h"
#include "cuda_util.h"

static_assert(sizeof(cudaIpcMemHandle_t) == B200PS_IPC_HANDLE_BYTES, "IPC handle size");

extern "C" int b200ps_dev_alloc(size_t bytes, void** d_ptr) {
    B200PS_REQUIRE(d_ptr && bytes > 0, B200PS_ERR_INVALID, "dev_alloc needs a destination and bytes > 0");
    B200PS_CUDA(cudaMalloc(d_ptr, bytes));
    return B200PS_OK;
}

extern "C" int b200ps_dev_free(void* d_ptr) {
    if (d_ptr) B200PS_CUDA(cudaFree(d_ptr));
    return B200PS_OK;
}

extern "C" int b200ps_ipc_export(const void* d_ptr, void* handle_out) {
    B200PS_REQUIRE(d_ptr && handle_out, B200PS_ERR_INVALID, "null argument");
    cudaIpcMemHandle_t h;
    B200PS_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
    memcpy(handle_out, &h, sizeof h);
    return B200PS_OK;
}

extern "C" int b200ps_ipc_open(const void* handle, void** d_ptr) {
    B200PS_REQUIRE(handle && d_ptr, B200PS_ERR_INVALID, "null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    B200PS_CUDA(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200PS_OK;
}

extern "C" int b200ps_ipc_close(void* d_ptr) {
    if (d_ptr) B200PS_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return B200PS_OK;
}

// ------------------------------------------------------------------------------------------------
// Stream-ordered barrier across the ranks of the box, through peer memory.
//
// Every rank owns a flag array (one 32-bit slot per rank) inside its exported arena.  Arriving at
// barrier number `epoch`, rank r stores `epoch` into slot r of EVERY rank's array (posted NVLink
// stores) and then waits until all slots of its OWN array have reached `epoch`.  The kernel that
// ran before it on the stream has completed, so its peer stores are already performed; the
// system-scope fences order the flag traffic itself.  One 32-thread CTA, ~ a few microseconds --
// against tens of microseconds for a library all-reduce used as a barrier.
//
// Every rank runs on its own GPU (one process per GPU), so the waiting kernels never wait for a
// kernel that needs the same GPU.  The wait is bounded: after ~4 s it gives up and raises the
// error word, which the host reads with b200ps_peer_barrier_error.
// ------------------------------------------------------------------------------------------------
namespace {

struct BarrierArgs {
    uint32_t* flags[B200PS_MAX_MIRRORS + 1];  // flags[r] = rank r's array (local or peer pointer)
    int rank, world;
    uint32_t epoch;
};

__global__ void peer_barrier_kernel(const __grid_constant__ BarrierArgs a) {
    const int t = threadIdx.x;
    if (t < a.world) {
        __threadfence_system();
        volatile uint32_t* remote = a.flags[t] + a.rank;
        *remote = a.epoch;  // my arrival, in rank t's array
        volatile uint32_t* mine = a.flags[a.rank] + t;
        const long long t0 = clock64();
        // epochs only grow; (int32) difference tolerates wrap-around
        while ((int32_t)(*mine - a.epoch) < 0) {
            if (clock64() - t0 > 8000000000ll) {  // ~4 s at 2 GHz: a peer died or never arrived
                a.flags[a.rank][B200PS_MAX_MIRRORS + 1] = 1u;
                break;
            }
        }
        __threadfence_system();
    }
}

}  // namespace

extern "C" int b200ps_peer_barrier(uint32_t* const* d_flags_by_rank, int rank, int world, uint32_t epoch, void* stream) {
    B200PS_REQUIRE(d_flags_by_rank, B200PS_ERR_INVALID, "null flag pointer list");
    B200PS_REQUIRE(world >= 1 && world <= B200PS_MAX_MIRRORS + 1 && rank >= 0 && rank < world, B200PS_ERR_RANGE,
                   "rank %d of world %d (at most %d ranks)", rank, world, B200PS_MAX_MIRRORS + 1);
    if (world == 1) return B200PS_OK;
    BarrierArgs a{};
    for (int r = 0; r < world; ++r) {
        B200PS_REQUIRE(d_flags_by_rank[r] && ((uintptr_t)d_flags_by_rank[r] & 3) == 0, B200PS_ERR_INVALID,
                       "flag array of rank %d is null or misaligned", r);
        a.flags[r] = d_flags_by_rank[r];
    }
    a.rank = rank;
    a.world = world;
    a.epoch = epoch;
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    B200PS_CUDA(cudaGetLastError());
    b200ps::g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_peer_barrier_error(const uint32_t* d_my_flags, void* stream) {
    B200PS_REQUIRE(d_my_flags, B200PS_ERR_INVALID, "null flag array");
    uint32_t e = 0;
    B200PS_CUDA(cudaMemcpyAsync(&e, d_my_flags + B200PS_MAX_MIRRORS + 1, sizeof e, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B200PS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B200PS_REQUIRE(e == 0, B200PS_ERR_STATE, "a peer barrier timed out: some rank never arrived");
    return B200PS_OK;
}

// ------------------------------------------------------------------------------------------------
// One-way signals through peer memory, for pipelining a step slice by slice: a rank that has
// finished producing slice s (its partial sum) stores the step number into the owner's flag word
// for (s, rank); the owner's stream waits for the words of all producing ranks before its apply
// kernel for s starts -- while the producers are already folding the next slices.
// ------------------------------------------------------------------------------------------------
namespace {

__global__ void peer_signal_kernel(uint32_t* flag, uint32_t epoch) {
    __threadfence_system();  // the producing kernel has completed; order the flag behind its stores
    *(volatile uint32_t*)flag = epoch;
}

__global__ void peer_wait_kernel(const uint32_t* flags, uint32_t rank_mask, uint32_t epoch, uint32_t* error_word) {
    const int t = threadIdx.x;
    if ((rank_mask >> t) & 1u) {
        const volatile uint32_t* f = flags + t;
        const long long t0 = clock64();
        while ((int32_t)(*f - epoch) < 0) {
            if (clock64() - t0 > 8000000000ll) {
                *error_word = 1u;
                break;
            }
        }
    }
    __threadfence_system();
}

}  // namespace

extern "C" int b200ps_peer_signal(uint32_t* d_flag, uint32_t epoch, void* stream) {
    B200PS_REQUIRE(d_flag && ((uintptr_t)d_flag & 3) == 0, B200PS_ERR_INVALID, "null or misaligned flag word");
    peer_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(d_flag, epoch);
    B200PS_CUDA(cudaGetLastError());
    b200ps::g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_peer_wait(const uint32_t* d_flags, uint32_t rank_mask, uint32_t epoch, uint32_t* d_error_word,
                                void* stream) {
    B200PS_REQUIRE(d_flags && d_error_word, B200PS_ERR_INVALID, "null flag array or error word");
    if (rank_mask == 0) return B200PS_OK;
    peer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(d_flags, rank_mask, epoch, d_error_word);
    B200PS_CUDA(cudaGetLastError());
    b200ps::g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}
