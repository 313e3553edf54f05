// peer_sync.cuh -- the stream-ordered barrier across the ranks of a box, through peer memory, whose
// epoch lives in device memory (so the launch is the same every time and a captured graph stays valid).
// Shared by the dense context (ctx.cu) and the embedding table (table.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "b200ps.h"

namespace b200ps {

constexpr int SYNC_MAX_RANKS = B200PS_MAX_MIRRORS + 1;
constexpr long long SYNC_SPIN_LIMIT = 8000000000ll;  // ~4 s at 2 GHz: a peer died or never arrived

struct PeerBarrierArgs {
    uint32_t* flags[SYNC_MAX_RANKS];  // flags[r] = rank r's barrier words (local or peer pointer)
    uint32_t* epoch;       // device word: barriers passed so far
    uint32_t* abort;       // device word: raised on a time-out; from then on barriers fall through
    uint32_t* host_abort;  // mapped pinned host word: lets the host see a time-out without synchronising
    uint32_t* step_epoch;  // optional device word advanced by a barrier that closes a pipelined step
    int rank, world;
};

// Arriving at its next barrier, rank r stores the new epoch into slot r of EVERY rank's flag array
// (posted NVLink stores) and waits until all slots of its OWN array have reached it.  The kernels
// queued before it on the stream have completed, so their peer stores are performed; the
// system-scope fences order the flag traffic itself.  Every rank runs on its own GPU, so a waiting
// kernel never waits for a kernel that needs the same GPU.
static __global__ void peer_barrier_epoch_kernel(const __grid_constant__ PeerBarrierArgs a) {
    __shared__ uint32_t epoch_sh;
    const int t = threadIdx.x;
    if (t == 0) {
        const uint32_t e = *a.epoch + 1;
        *a.epoch = e;
        epoch_sh = e;
        if (a.step_epoch) *a.step_epoch += 1;
    }
    __syncthreads();
    const uint32_t epoch = epoch_sh;
    if (t < a.world && !*(volatile uint32_t*)a.abort) {
        __threadfence_system();
        *(volatile uint32_t*)(a.flags[t] + a.rank) = epoch;  // my arrival, in rank t's array
        const volatile uint32_t* mine = a.flags[a.rank] + t;
        const long long t0 = clock64();
        while ((int32_t)(*mine - epoch) < 0) {  // epochs only grow; the signed difference tolerates wrap-around
            if (clock64() - t0 > SYNC_SPIN_LIMIT) {
                *(volatile uint32_t*)a.abort = 1u;
                *(volatile uint32_t*)a.host_abort = 1u;
                break;
            }
        }
        __threadfence_system();
    }
}

}  // namespace b200ps
