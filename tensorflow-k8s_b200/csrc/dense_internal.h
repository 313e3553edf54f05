// dense_internal.h -- what ctx.cu needs from dense.cu beyond the public stateless entry points:
// a launch with a per-context geometry and the context's abort word, and the hooks graph replay
// needs to change a captured launch's step size.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "b200ps.h"

namespace b200ps {

// FOLD: out[i] = left fold of the buffers, nothing else
enum { OPT_SGD = 0, OPT_ADAM = 1, OPT_FOLD = 2, OPT_MOMENTUM = 3, OPT_FTRL = 4 };

struct Tuning {
    // measured on B200 (profiles/r1_dense_sweep.jsonl): 256 x 8 x 2 streams Adam W=4 at 6.5 TB/s
    int threads = 256;
    int ctas_per_sm = 8;
    int unroll = 2;
};

struct DenseLaunch {
    int opt = OPT_SGD;
    float* p = nullptr;        // slice (FOLD: destination)
    float* s1 = nullptr;       // Adam m | momentum accumulator | FTRL accum
    float* s2 = nullptr;       // Adam v | FTRL linear
    const float* const* grads = nullptr;  // host array of W device pointers, fold order
    int W = 0;
    float* const* mirrors = nullptr;      // host array of n_mirrors device (peer) pointers
    int n_mirrors = 0;
    uint64_t n = 0;
    float inv_w = 1.f;
    float lr = 0.f;            // SGD / momentum / FTRL: lr.  Adam: lr_t.
    float c1 = 0.f, c2 = 0.f;  // Adam beta1, beta2 | momentum, - | FTRL l1, 2*l2
    float omb1 = 0.f, omb2 = 0.f, eps = 0.f;
    const uint32_t* abort_word = nullptr;
    // pipelined step: wait in the kernel's prologue until flags[r] >= *wait_epoch + 1 for every r in wait_mask
    const uint32_t* wait_flags = nullptr;
    const uint32_t* wait_epoch = nullptr;
    uint32_t wait_mask = 0;
    uint32_t* abort_out = nullptr;
};

// Apply of one slice in B200PS_AGG_MULTIMEM mode: the gradient sum comes from the multicast mapping of
// the partial-sum buffers (added in the switch), the updated slice goes to every replica through the
// multicast mapping of the replicas.  (Field names match DenseParams: the optimizer arithmetic is shared.)
struct McApply {
    int opt = OPT_SGD;
    const float* g_mc = nullptr;  // multicast address of this slice in the partial-sum buffers
    float* p_mc = nullptr;        // multicast address of this slice in the replicas
    const float* p = nullptr;     // this GPU's replica (read)
    float* m = nullptr;
    float* v = nullptr;
    uint64_t n = 0;
    uint64_t rot4 = 0;            // the walk over the slice starts here (float4 units) and wraps around
    float inv_w = 1.f, lr = 0.f, beta1 = 0.f, beta2 = 0.f, omb1 = 0.f, omb2 = 0.f, eps = 0.f;
    const uint32_t* abort_word = nullptr;
    const uint32_t* wait_flags = nullptr;
    const uint32_t* wait_epoch = nullptr;
    uint32_t wait_mask = 0;
    uint32_t* abort_out = nullptr;
};
int dense_launch_mc(const McApply& a, int ctas_per_sm, cudaStream_t stream);

// One kernel folds every slice of the flat buffer in order and publishes each to its owner (see dense.cu).
struct FoldPublish {
    const float* g[B200PS_MAX_WORKERS];  // the local compute tasks' gradient buffers (whole flat buffer), worker order
    int W;
    float* out;                          // this GPU's partial-sum buffer (whole flat buffer)
    int n_slices;
    uint64_t begin[64], end[64];         // slice ranges (elements)
    uint32_t* flag[64];                  // where slice s is published: the owner's flag word for this rank (local or peer)
    uint32_t* done;                      // n_slices device counters, zero between steps
    uint32_t* epoch;                     // device word: this step publishes *epoch + 1 and then stores it back
    const uint32_t* abort_word;
};
int dense_fold_publish(const FoldPublish& f, int ctas_per_sm, cudaStream_t stream);

int dense_launch(const DenseLaunch& l, const Tuning& t, cudaStream_t stream);

// Graph replay: OPT_* of a graph node if it is one of dense.cu's kernels (else -1), and a new step
// size for such a node in an instantiated graph (affects launches enqueued afterwards).
int dense_node_opt(cudaGraphNode_t node);
cudaError_t dense_node_set_lr(cudaGraphExec_t exec, cudaGraphNode_t node, float lr);

}  // namespace b200ps
