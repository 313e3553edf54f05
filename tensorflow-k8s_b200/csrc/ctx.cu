// ctx.cu -- the PS tasks and compute-task buffers of one GPU as one object the library owns
// (include/b200ps.h "Context group"; SURVEY.md 8(b) dense group, 8(e) multi-GPU).
//
// Reference: none -- what TensorFlow's PS runtime does inside the containers a TFJob launches
// (variables and optimizer slots on the PS tasks, gRPC transport to the WORKER tasks;
// k8s-operator.md:6).  Here one process per GPU owns an arena in HBM that the other ranks of the
// box map (CUDA IPC), and a step is: [local fold] -> barrier -> one fused kernel per local shard
// (peer loads = push, fixed-order fold, optimizer, peer stores = pull) -> barrier.  Barriers are
// flag exchanges through peer memory whose epoch lives in device memory, so the whole step of a
// rank is a fixed sequence of launches that can be replayed as ONE CUDA graph; the only thing that
// changes from step to step -- Adam's bias-corrected step size -- stays a kernel parameter (reading
// it from device memory cost 4 % of the step) and is re-armed on the graph's apply nodes before
// each replay.
#include <cuda_runtime.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "common.h"
#include "cuda_util.h"
#include "dense_internal.h"
#include "peer_sync.cuh"
#include "plan.h"
#include "vmm.h"

namespace b200ps {
namespace {

constexpr int MAX_RANKS = SYNC_MAX_RANKS;
constexpr int SLICE_FLAG_STRIDE = 16;  // words per slice in the slice-ready flags: one per producing rank
constexpr int MAX_SLICES = 64;
constexpr int FLAG_WORDS = B200PS_BARRIER_WORDS + MAX_SLICES * SLICE_FLAG_STRIDE;  // behind every arena

// What a kernel needs that changes from step to step, in device memory, so that a captured graph
// stays valid: the barrier epoch and the abort word.
struct DevState {
    uint32_t epoch;       // barrier epoch
    uint32_t abort;
    uint32_t step_epoch;  // pipelined steps completed (advanced by the step's closing barrier): what the apply kernels wait for, minus 1
    uint32_t fold_epoch;  // pipelined folds completed (advanced by the fold kernel): what it publishes, minus 1
    uint32_t done[MAX_SLICES];  // arrival counters of the publishing fold kernel
};

const char* opt_name(int o) {
    return o == B200PS_OPT_ADAM ? "adam" : o == B200PS_OPT_MOMENTUM ? "momentum" : o == B200PS_OPT_FTRL ? "ftrl" : "sgd";
}

}  // namespace
}  // namespace b200ps

using namespace b200ps;

struct b200ps_ctx {
    std::unique_ptr<b200ps_plan> plan;
    b200ps_dense_options_t o{};
    Tuning tune;
    int device = 0;
    float inv_w = 1.f;
    uint64_t t = 0;            // optimizer steps taken (Adam's t)
    float* arena = nullptr;    // own allocation (arena_elems + FLAG_WORDS words)
    float* base[MAX_RANKS] = {nullptr};  // base[r] = rank r's arena as mapped here
    bool imported = false;
    // B200PS_AGG_MULTIMEM: the arena is a VMM allocation, replica + partial sum are bound to a multicast object
    bool multimem = false;
    vmm::Block vm_own;
    vmm::Block vm_peer[MAX_RANKS];
    vmm::Multicast mc;
    std::unique_ptr<vmm::FdServer> fd_server;
    std::string peer_server[MAX_RANKS];
    size_t vm_bytes = 0, mc_bytes = 0;
    std::vector<float*> s1, s2;  // per launch: optimizer state of the local shard
    DevState* st = nullptr;
    uint32_t* host_abort = nullptr;      // pinned, mapped
    uint32_t* host_abort_dev = nullptr;  // its device alias
    cudaStream_t s_in = nullptr, s_out = nullptr, s_fold = nullptr;
    cudaEvent_t ev_fold = nullptr;
    bool pipelined = false;  // two-level folds published slice by slice, applies waiting in their prologue
    std::vector<cudaEvent_t> ev_in, ev_k, ev_out;  // per chunk of step_host
    std::vector<char> ev_out_valid;
    cudaEvent_t ev_join = nullptr, ev_d2h = nullptr, ev_fork = nullptr;
    bool out_busy = false;  // a copy-out of step_host may still be in flight on s_out
    b200ps_adam_t adam_c{};  // Adam's constants (lr_t is evaluated per step)
    // graph replay
    cudaGraphExec_t graph = nullptr;
    struct GraphDel { void operator()(CUgraph_st* g) const { if (g) cudaGraphDestroy(g); } };
    std::unique_ptr<CUgraph_st, GraphDel> graph_src;  // kept alive: the node handles below belong to it
    uint64_t graph_kernels = 0;  // kernel nodes of the graph (b200ps_launch_count counts them per replay)
    std::vector<cudaGraphNode_t> graph_adam_nodes;  // the apply nodes whose step size is re-armed per replay
    // profiling
    bool profile = false;
    struct Span { cudaEvent_t a, b; };
    std::vector<Span> spans[3];  // fold, apply, barrier
    std::vector<cudaEvent_t> event_pool;
    std::vector<cudaEvent_t> phase_ev;
    int phase_ev_mode = 0;

    uint32_t* flags_of(int r) const { return reinterpret_cast<uint32_t*>(base[r] + plan->arena_elems(r)); }
    bool adam() const { return o.optimizer == B200PS_OPT_ADAM; }
};

namespace {

int check_alive(const b200ps_ctx* c) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    B200PS_REQUIRE(*(volatile uint32_t*)c->host_abort == 0, B200PS_ERR_STATE,
                   "a peer barrier of this context timed out (some rank never arrived); the parameters were not updated since");
    return B200PS_OK;
}

int need_peers(b200ps_ctx* c) {
    B200PS_REQUIRE(c->plan->world == 1 || c->imported, B200PS_ERR_STATE,
                   "world = %d: call b200ps_ctx_import_handles before using the context", c->plan->world);
    if (c->multimem && !c->mc.bound) return vmm::mc_bind_and_map(&c->mc, c->device, c->vm_own);
    return B200PS_OK;
}

float lr_for_step(const b200ps_ctx* c, uint64_t t) {
    if (!c->adam()) return (float)c->o.lr;
    b200ps_adam_t a;
    b200ps_adam_scalars(c->o.lr, c->o.beta1, c->o.beta2, c->o.eps, t, &a);
    return a.lr_t;
}

// a step_host copy-out may still be reading the replica: order `st` behind it
int wait_copy_out(b200ps_ctx* c, cudaStream_t st) {
    if (!c->out_busy) return B200PS_OK;
    B200PS_CUDA(cudaEventRecord(c->ev_d2h, c->s_out));
    B200PS_CUDA(cudaStreamWaitEvent(st, c->ev_d2h, 0));
    c->out_busy = false;
    for (auto& v : c->ev_out_valid) v = 0;
    return B200PS_OK;
}

void drop_graph(b200ps_ctx* c) {
    if (c->graph) cudaGraphExecDestroy(c->graph);
    c->graph = nullptr;
    c->graph_adam_nodes.clear();
    c->graph_src.reset();
}

cudaEvent_t pooled_event(b200ps_ctx* c) {
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    c->event_pool.push_back(e);
    return e;
}

struct Mark {
    b200ps_ctx* c;
    int phase;
    cudaStream_t st;
    cudaEvent_t a = nullptr;
    Mark(b200ps_ctx* c_, int phase_, cudaStream_t st_) : c(c_), phase(phase_), st(st_) {
        if (c->profile && (a = pooled_event(c))) cudaEventRecord(a, st);
    }
    ~Mark() {
        if (a) {
            cudaEvent_t b = pooled_event(c);
            if (b) {
                cudaEventRecord(b, st);
                c->spans[phase].push_back({a, b});
            }
        }
    }
};

int barrier(b200ps_ctx* c, cudaStream_t st, bool advance_step = false) {
    const auto& p = *c->plan;
    if (p.world == 1) return B200PS_OK;
    Mark m(c, 2, st);
    PeerBarrierArgs a{};
    for (int r = 0; r < p.world; ++r) a.flags[r] = c->flags_of(r);
    a.epoch = &c->st->epoch;
    a.abort = &c->st->abort;
    a.host_abort = c->host_abort_dev;
    a.step_epoch = advance_step ? &c->st->step_epoch : nullptr;
    a.rank = p.rank;
    a.world = p.world;
    peer_barrier_epoch_kernel<<<1, 32, 0, st>>>(a);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

// One kernel over elements [lo, hi) of launch i's slice (flat-buffer coordinates).  `only_src` >= 0
// folds just that one source (ordered-async mode); mirrors can be suppressed.
int launch_apply(b200ps_ctx* c, size_t i, uint64_t lo, uint64_t hi, uint64_t t_step, cudaStream_t st, int only_src = -1,
                 bool mirrors = true, float inv_w_override = 0.f, int wait_unit = -1) {
    const auto& p = *c->plan;
    const auto& L = p.launches[i];
    if (hi <= lo) return B200PS_OK;
    const float* g[B200PS_MAX_WORKERS];
    float* mir[B200PS_MAX_MIRRORS];
    int ng = 0, nm = 0;
    for (size_t k = 0; k < L.src.size(); ++k) {
        if (only_src >= 0 && (int)k != only_src) continue;
        g[ng++] = c->base[L.src[k].first] + L.src[k].second + lo;
    }
    if (mirrors)
        for (int r : L.mirrors) mir[nm++] = c->base[r] + lo;
    if (c->multimem && only_src < 0) {
        McApply a;
        float* mc = reinterpret_cast<float*>(c->mc.va);
        a.g_mc = mc + p.P + lo;  // the partial-sum buffers sit behind the replicas in every arena
        a.p_mc = mc + lo;
        a.p = c->arena + lo;
        a.n = hi - lo;
        if (const char* rot = getenv("B200PS_MC_ROTATE"))  // experiment knob: rank-dependent start of the walk
            if (atoi(rot)) a.rot4 = ((uint64_t)p.rank * (a.n / 4) / (uint64_t)p.world) & ~7ull;
        a.inv_w = c->inv_w;
        a.abort_word = &c->st->abort;
        if (wait_unit >= 0) {
            a.wait_flags = c->flags_of(p.rank) + B200PS_BARRIER_WORDS + wait_unit * SLICE_FLAG_STRIDE;
            a.wait_epoch = &c->st->step_epoch;
            a.abort_out = &c->st->abort;
            a.wait_mask = (1u << p.world) - 1u;
        }
        const uint64_t so = lo - L.begin;
        a.m = c->s1[i] ? c->s1[i] + so : nullptr;
        a.v = c->s2[i] ? c->s2[i] + so : nullptr;
        switch (c->o.optimizer) {
            case B200PS_OPT_ADAM:
                a.opt = OPT_ADAM;
                a.lr = lr_for_step(c, t_step);
                a.beta1 = c->adam_c.beta1;
                a.beta2 = c->adam_c.beta2;
                a.omb1 = c->adam_c.omb1;
                a.omb2 = c->adam_c.omb2;
                a.eps = c->adam_c.eps;
                break;
            case B200PS_OPT_MOMENTUM:
                a.opt = OPT_MOMENTUM;
                a.lr = (float)c->o.lr;
                a.beta1 = (float)c->o.momentum;
                break;
            case B200PS_OPT_FTRL:
                a.opt = OPT_FTRL;
                a.lr = (float)c->o.lr;
                a.beta1 = (float)c->o.ftrl_l1;
                a.beta2 = 2.0f * (float)c->o.ftrl_l2;
                break;
            default:
                a.opt = OPT_SGD;
                a.lr = (float)c->o.lr;
        }
        return dense_launch_mc(a, wait_unit >= 0 ? 2 : 8, st);
    }
    DenseLaunch d;
    d.p = c->arena + lo;
    d.grads = g;
    d.W = ng;
    d.mirrors = mir;
    d.n_mirrors = nm;
    d.n = hi - lo;
    d.inv_w = inv_w_override != 0.f ? inv_w_override : c->inv_w;
    d.abort_word = &c->st->abort;
    if (wait_unit >= 0) {  // the sources are published by the producing ranks' fold kernels (flags in this rank's arena)
        d.wait_flags = c->flags_of(p.rank) + B200PS_BARRIER_WORDS + wait_unit * SLICE_FLAG_STRIDE;
        d.wait_epoch = &c->st->step_epoch;
        d.abort_out = &c->st->abort;
        for (const auto& src : L.src) d.wait_mask |= 1u << src.first;
    }
    const uint64_t off = lo - L.begin;
    switch (c->o.optimizer) {
        case B200PS_OPT_ADAM: {
            const b200ps_adam_t& a = c->adam_c;
            d.opt = OPT_ADAM;
            d.s1 = c->s1[i] + off;
            d.s2 = c->s2[i] + off;
            d.c1 = a.beta1;
            d.c2 = a.beta2;
            d.omb1 = a.omb1;
            d.omb2 = a.omb2;
            d.eps = a.eps;
            d.lr = lr_for_step(c, t_step);  // lr * sqrt(1 - beta2^t) / (1 - beta1^t), evaluated in double
            break;
        }
        case B200PS_OPT_MOMENTUM:
            d.opt = OPT_MOMENTUM;
            d.s1 = c->s1[i] + off;
            d.lr = (float)c->o.lr;
            d.c1 = (float)c->o.momentum;
            break;
        case B200PS_OPT_FTRL:
            d.opt = OPT_FTRL;
            d.s1 = c->s1[i] + off;
            d.s2 = c->s2[i] + off;
            d.lr = (float)c->o.lr;
            d.c1 = (float)c->o.ftrl_l1;
            d.c2 = 2.0f * (float)c->o.ftrl_l2;
            break;
        default:
            d.opt = OPT_SGD;
            d.lr = (float)c->o.lr;
    }
    return dense_launch(d, c->tune, st);
}

// local half of the two-level aggregation over [lo, hi): partial = fold of this GPU's workers
int launch_fold(b200ps_ctx* c, uint64_t lo, uint64_t hi, cudaStream_t st) {
    const auto& p = *c->plan;
    if (!p.two_level() || p.local_workers.empty() || hi <= lo) return B200PS_OK;
    const float* g[B200PS_MAX_WORKERS];
    int ng = 0;
    for (int w : p.local_workers) g[ng++] = c->arena + p.grad_offset(w).second + lo;
    DenseLaunch d;
    d.opt = OPT_FOLD;
    d.p = c->arena + p.P + lo;
    d.grads = g;
    d.W = ng;
    d.n = hi - lo;
    d.abort_word = &c->st->abort;
    return dense_launch(d, c->tune, st);
}

// Pipelined form of the step (G > 1, two-level): the fold kernel runs on its own stream and publishes
// slice after slice; the apply kernels, launched at once on `st` with a smaller grid, wait in their
// prologue for the slices they own.  One barrier closes the step: every replica holds every updated
// slice, and every partial sum may be overwritten.
int enqueue_step_pipelined(b200ps_ctx* c, cudaStream_t st) {
    const auto& p = *c->plan;
    // publish units: every slice is cut into C chunks, walked chunk-major (chunk 0 of every slice, then
    // chunk 1, ...), so that every owner gets work early; at most 32 units keep the per-unit
    // synchronisation (a fence and one atomic per CTA) small against the fold itself
    const int C = p.S >= 32 ? 1 : (32 / p.S < 4 ? 32 / p.S : 4);
    auto chunk_range = [&](int s, int k, uint64_t* lo, uint64_t* hi) {
        const uint64_t b = p.slices[(size_t)s].first, e = p.slices[(size_t)s].second;
        const uint64_t len = ((e - b + (uint64_t)C - 1) / (uint64_t)C + 1023) / 1024 * 1024;
        *lo = b + (uint64_t)k * len < e ? b + (uint64_t)k * len : e;
        *hi = *lo + len < e ? *lo + len : e;
    };
    B200PS_CUDA(cudaEventRecord(c->ev_fork, st));
    B200PS_CUDA(cudaStreamWaitEvent(c->s_fold, c->ev_fork, 0));
    if (!p.local_workers.empty()) {
        FoldPublish f{};
        for (int w : p.local_workers) f.g[f.W++] = c->arena + p.grad_offset(w).second;
        f.out = c->arena + p.P;
        for (int k = 0; k < C; ++k)
            for (int s = 0; s < p.S; ++s) {
                const int u = f.n_slices++;
                chunk_range(s, k, &f.begin[u], &f.end[u]);
                const int owner = b200ps_placement_gpu(s, p.world);
                f.flag[u] = c->flags_of(owner) + B200PS_BARRIER_WORDS + u * SLICE_FLAG_STRIDE + p.rank;
            }
        f.done = c->st->done;
        f.epoch = &c->st->fold_epoch;
        f.abort_word = &c->st->abort;
        if (int rc = dense_fold_publish(f, 2, c->s_fold)) return rc;
    }
    B200PS_CUDA(cudaEventRecord(c->ev_fold, c->s_fold));
    // Both kernels must be resident together whatever order the streams start them in: per SM two fold CTAs
    // (256 threads x 55 registers) and two apply CTAs (256 x <= 64 registers at unroll 1) fit the register file.
    // The apply is NVLink-bound, so the small grid costs nothing.
    Tuning keep = c->tune;
    c->tune = Tuning{256, 2, 1};
    int rc = B200PS_OK;
    for (int k = 0; k < C && !rc; ++k)
        for (size_t i = 0; i < p.launches.size() && !rc; ++i) {
            uint64_t lo = 0, hi = 0;
            chunk_range(p.launches[i].shard, k, &lo, &hi);
            rc = launch_apply(c, i, lo, hi, c->t + 1, st, -1, true, 0.f, k * p.S + p.launches[i].shard);
        }
    c->tune = keep;
    if (rc) return rc;
    B200PS_CUDA(cudaStreamWaitEvent(st, c->ev_fold, 0));
    return barrier(c, st, true);
}

// experiment knob (B200PS_STEP_EVENTS = 1 | 2 | 3): event records behind every phase of the sequential step
// (1: without timing, 2: with timing, 3: two timing events back to back, as the profile pass leaves them)
void phase_event(b200ps_ctx* c, cudaStream_t st, int k) {
    const char* env = getenv("B200PS_STEP_EVENTS");
    const int mode = env ? atoi(env) : 0;
    if (!mode || c->profile) return;
    if (c->phase_ev.size() < 8 || c->phase_ev_mode != mode) {
        for (auto& e : c->phase_ev)
            if (e) cudaEventDestroy(e);
        c->phase_ev.assign(8, nullptr);
        for (auto& e : c->phase_ev) cudaEventCreateWithFlags(&e, mode == 1 ? cudaEventDisableTiming : cudaEventDefault);
        c->phase_ev_mode = mode;
    }
    cudaEventRecord(c->phase_ev[(size_t)k], st);
    if (mode == 3) cudaEventRecord(c->phase_ev[(size_t)k + 4], st);
}

int enqueue_step(b200ps_ctx* c, cudaStream_t st) {
    const auto& p = *c->plan;
    if (c->pipelined && !c->profile) return enqueue_step_pipelined(c, st);
    {
        Mark m(c, 0, st);
        if (int rc = launch_fold(c, 0, p.P, st)) return rc;
    }
    phase_event(c, st, 0);
    if (int rc = barrier(c, st)) return rc;  // every worker's gradient (or every GPU's partial sum) is in place
    phase_event(c, st, 1);
    {
        Mark m(c, 1, st);
        for (size_t i = 0; i < p.launches.size(); ++i)
            if (int rc = launch_apply(c, i, p.launches[i].begin, p.launches[i].end, c->t + 1, st)) return rc;
    }
    phase_event(c, st, 2);
    const int rc = barrier(c, st);  // every replica holds every updated slice
    phase_event(c, st, 3);
    return rc;
}

int zero_state(b200ps_ctx* c, cudaStream_t st) {
    const auto& p = *c->plan;
    for (size_t i = 0; i < p.launches.size(); ++i) {
        const size_t n = (size_t)(p.launches[i].end - p.launches[i].begin);
        if (c->s1[i]) {
            if (c->o.optimizer == B200PS_OPT_FTRL) {  // TF's initial_accumulator_value
                std::vector<float> init(n, (float)c->o.ftrl_accum_init);
                B200PS_CUDA(cudaMemcpyAsync(c->s1[i], init.data(), n * 4, cudaMemcpyHostToDevice, st));
            } else {
                B200PS_CUDA(cudaMemsetAsync(c->s1[i], 0, n * 4, st));
            }
        }
        if (c->s2[i]) B200PS_CUDA(cudaMemsetAsync(c->s2[i], 0, n * 4, st));
    }
    c->t = 0;
    return B200PS_OK;
}

std::string ckpt_path(const b200ps_ctx* c, const char* dir, int r) {
    return std::string(dir) + "/ps-rank" + std::to_string(r) + "-of-" + std::to_string(c->plan->world) + ".b200ps";
}

struct CkptHeader {
    char magic[8];  // "B200PSD2"
    uint64_t P, t;
    int32_t S, W, optimizer, rank, world, n_shards;
};
struct CkptShard {
    int32_t shard, n_state;
    uint64_t begin, end;
};

}  // namespace

// ------------------------------------------------------------------------------------------------
extern "C" int b200ps_dense_options_default(b200ps_dense_options_t* o) {
    B200PS_REQUIRE(o, B200PS_ERR_INVALID, "null argument");
    memset(o, 0, sizeof *o);
    o->optimizer = B200PS_OPT_SGD;
    o->aggregate = B200PS_AGG_TWO_LEVEL;
    o->lr = 0.01;
    o->beta1 = 0.9;
    o->beta2 = 0.999;
    o->eps = 1e-8;
    o->momentum = 0.9;
    o->ftrl_accum_init = 0.1;
    return B200PS_OK;
}

extern "C" int b200ps_ctx_create(const b200ps_cluster_t* cluster, const b200ps_shardmap_t* map, int rank, int world,
                                 const b200ps_dense_options_t* opt, b200ps_ctx_t** out) {
    B200PS_REQUIRE(cluster && map && out, B200PS_ERR_INVALID, "null argument");
    b200ps_dense_options_t o;
    b200ps_dense_options_default(&o);
    if (opt) o = *opt;
    B200PS_REQUIRE(o.optimizer >= B200PS_OPT_SGD && o.optimizer <= B200PS_OPT_FTRL, B200PS_ERR_INVALID, "unknown optimizer %d",
                   o.optimizer);
    B200PS_REQUIRE(o.lr > 0, B200PS_ERR_RANGE, "lr must be positive");
    const int S = b200ps_cluster_num_shards(cluster), W = b200ps_cluster_num_workers(cluster);
    B200PS_REQUIRE(S == b200ps_shard_map_nshards(map), B200PS_ERR_INVALID,
                   "the cluster has %d ps tasks but the shard map has %d shards", S, b200ps_shard_map_nshards(map));
    auto c = std::make_unique<b200ps_ctx>();
    b200ps_plan_t* plan = nullptr;
    if (int rc = b200ps_plan_build(map, W, rank, world, o.aggregate, &plan)) return rc;
    c->plan.reset(plan);
    c->o = o;
    if (o.threads || o.ctas_per_sm || o.unroll) {
        B200PS_REQUIRE((!o.threads || (o.threads >= 32 && o.threads <= 1024 && o.threads % 32 == 0)) &&
                           (!o.ctas_per_sm || (o.ctas_per_sm >= 1 && o.ctas_per_sm <= 32)) &&
                           (!o.unroll || o.unroll == 1 || o.unroll == 2 || o.unroll == 4),
                       B200PS_ERR_RANGE, "launch geometry: threads a multiple of 32 in 32..1024, ctas_per_sm 1..32, unroll 1|2|4");
        if (o.threads) c->tune.threads = o.threads;
        if (o.ctas_per_sm) c->tune.ctas_per_sm = o.ctas_per_sm;
        if (o.unroll) c->tune.unroll = o.unroll;
    }
    if (!c->o.host_chunk_elems) c->o.host_chunk_elems = 4u << 20;
    c->o.host_chunk_elems = c->o.host_chunk_elems / 1024 * 1024;
    if (c->o.host_chunk_elems < 1024) c->o.host_chunk_elems = 1024;
    c->inv_w = 1.0f / (float)W;  // one fp32 division, as the oracle forms it
    b200ps_adam_scalars(o.lr, o.beta1, o.beta2, o.eps, 1, &c->adam_c);
    B200PS_CUDA(cudaGetDevice(&c->device));
    const size_t words = (size_t)plan->arena_elems(rank) + FLAG_WORDS;
    B200PS_REQUIRE(words > FLAG_WORDS || plan->P == 0, B200PS_ERR_RANGE, "empty arena");
    c->multimem = o.aggregate == B200PS_AGG_MULTIMEM && plan->two_level();
    if (c->multimem) {
        B200PS_CUDA(cudaFree(nullptr));  // the runtime's context is current before the driver calls
        B200PS_REQUIRE(vmm::available(), B200PS_ERR_STATE,
                       "B200PS_AGG_MULTIMEM needs NVSwitch multicast support (b200ps_multicast_supported() is 0 on this device)");
        for (int r = 0; r < world; ++r)
            B200PS_REQUIRE(plan->workers_on(r) > 0, B200PS_ERR_INVALID,
                           "B200PS_AGG_MULTIMEM adds the partial sums of ALL GPUs in the switch: every GPU must host a compute task");
        const size_t gran = vmm::granularity(c->device, (size_t)plan->P * 8, world);
        c->mc_bytes = ((size_t)plan->P * 8 + gran - 1) / gran * gran;  // replica + partial sum
        c->vm_bytes = (words * 4 + gran - 1) / gran * gran;
        if (c->vm_bytes < c->mc_bytes) c->vm_bytes = c->mc_bytes;
        if (int rc = vmm::alloc(c->device, c->vm_bytes, &c->vm_own)) return rc;
        c->arena = reinterpret_cast<float*>(c->vm_own.va);
        B200PS_CUDA(cudaMemset(c->arena, 0, c->vm_bytes));
    } else {
        B200PS_CUDA(cudaMalloc(&c->arena, words * 4));
    }
    c->base[rank] = c->arena;
    B200PS_CUDA(cudaMemset(c->arena + plan->arena_elems(rank), 0, FLAG_WORDS * 4));
    B200PS_CUDA(cudaMalloc(&c->st, sizeof(DevState)));
    B200PS_CUDA(cudaMemset(c->st, 0, sizeof(DevState)));
    B200PS_CUDA(cudaHostAlloc(&c->host_abort, sizeof(uint32_t), cudaHostAllocMapped));
    *c->host_abort = 0;
    B200PS_CUDA(cudaHostGetDevicePointer(&c->host_abort_dev, c->host_abort, 0));
    const int n_state = o.optimizer == B200PS_OPT_SGD ? 0 : o.optimizer == B200PS_OPT_MOMENTUM ? 1 : 2;
    for (const auto& L : plan->launches) {
        const size_t n = (size_t)(L.end - L.begin);
        float *a = nullptr, *b = nullptr;
        if (n_state >= 1 && n) B200PS_CUDA(cudaMalloc(&a, n * 4));
        if (n_state >= 2 && n) B200PS_CUDA(cudaMalloc(&b, n * 4));
        c->s1.push_back(a);
        c->s2.push_back(b);
    }
    B200PS_CUDA(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    B200PS_CUDA(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    B200PS_CUDA(cudaStreamCreateWithFlags(&c->s_fold, cudaStreamNonBlocking));
    B200PS_CUDA(cudaEventCreateWithFlags(&c->ev_fold, cudaEventDisableTiming));
    c->pipelined = o.pipeline_folds && plan->two_level() && plan->S <= MAX_SLICES && plan->world <= SLICE_FLAG_STRIDE;
    B200PS_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    B200PS_CUDA(cudaEventCreateWithFlags(&c->ev_d2h, cudaEventDisableTiming));
    B200PS_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    if (int rc = zero_state(c.get(), nullptr)) return rc;
    B200PS_CUDA(cudaDeviceSynchronize());
    *out = c.release();
    return B200PS_OK;
}

extern "C" int b200ps_ctx_destroy(b200ps_ctx_t* c) {
    if (!c) return B200PS_OK;
    cudaDeviceSynchronize();
    drop_graph(c);
    if (c->multimem) {
        if (c->fd_server) c->fd_server->stop();
        vmm::mc_release(&c->mc);
        for (int r = 0; r < MAX_RANKS; ++r) vmm::release(&c->vm_peer[r]);
    } else {
        for (int r = 0; r < MAX_RANKS; ++r)
            if (c->base[r] && c->base[r] != c->arena) cudaIpcCloseMemHandle(c->base[r]);
    }
    for (float* p : c->s1) cudaFree(p);
    for (float* p : c->s2) cudaFree(p);
    for (auto& v : {c->ev_in, c->ev_k, c->ev_out})
        for (cudaEvent_t e : v) cudaEventDestroy(e);
    for (cudaEvent_t e : c->event_pool) cudaEventDestroy(e);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev_d2h) cudaEventDestroy(c->ev_d2h);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    if (c->s_fold) cudaStreamDestroy(c->s_fold);
    if (c->ev_fold) cudaEventDestroy(c->ev_fold);
    cudaFree(c->st);
    cudaFreeHost(c->host_abort);
    if (c->multimem)
        vmm::release(&c->vm_own);
    else
        cudaFree(c->arena);
    delete c;
    return B200PS_OK;
}

namespace {

// the 64-byte handle of a VMM arena: where its file descriptor can be fetched (csrc/vmm.cu)
struct VmHandle {
    char magic[4];  // "B2VM"
    int32_t rank;
    uint64_t bytes, mc_bytes;
    char server[40];
};
static_assert(sizeof(VmHandle) <= B200PS_IPC_HANDLE_BYTES, "handle size");

}  // namespace

extern "C" int b200ps_ctx_export_handle(b200ps_ctx_t* c, void* handle_out) {
    B200PS_REQUIRE(c && handle_out, B200PS_ERR_INVALID, "null argument");
    if (!c->multimem) return b200ps_ipc_export(c->arena, handle_out);
    B200PS_REQUIRE(!c->fd_server, B200PS_ERR_STATE, "the handle was already exported");
    int arena_fd = -1, mc_fd = -1;
    if (int rc = vmm::export_fd(c->vm_own, &arena_fd)) return rc;
    if (c->plan->rank == 0)  // rank 0 creates the multicast object and serves its handle too
        if (int rc = vmm::mc_create(c->plan->world, c->mc_bytes, &c->mc, &mc_fd)) return rc;
    static std::atomic<int> counter{0};
    char name[40];
    snprintf(name, sizeof name, "b200ps-%ld-%d", (long)getpid(), counter.fetch_add(1));
    c->fd_server = std::make_unique<vmm::FdServer>();
    if (int rc = c->fd_server->start(name, arena_fd, mc_fd)) return rc;
    VmHandle h{};
    memcpy(h.magic, "B2VM", 4);
    h.rank = c->plan->rank;
    h.bytes = c->vm_bytes;
    h.mc_bytes = c->mc_bytes;
    snprintf(h.server, sizeof h.server, "%s", name);
    memset(handle_out, 0, B200PS_IPC_HANDLE_BYTES);
    memcpy(handle_out, &h, sizeof h);
    return B200PS_OK;
}

extern "C" int b200ps_ctx_import_handles(b200ps_ctx_t* c, const void* handles) {
    B200PS_REQUIRE(c && (handles || c->plan->world == 1), B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(!c->imported, B200PS_ERR_STATE, "handles were already imported");
    const int world = c->plan->world, rank = c->plan->rank;
    if (!c->multimem) {
        for (int r = 0; r < world; ++r) {
            if (r == rank) continue;
            void* p = nullptr;
            if (int rc = b200ps_ipc_open((const char*)handles + (size_t)r * B200PS_IPC_HANDLE_BYTES, &p)) return rc;
            c->base[r] = (float*)p;
        }
        c->imported = true;
        return B200PS_OK;
    }
    for (int r = 0; r < world; ++r) {
        VmHandle h;
        memcpy(&h, (const char*)handles + (size_t)r * B200PS_IPC_HANDLE_BYTES, sizeof h);
        B200PS_REQUIRE(memcmp(h.magic, "B2VM", 4) == 0 && h.rank == r && h.mc_bytes == c->mc_bytes, B200PS_ERR_INVALID,
                       "handle %d is not the B200PS_AGG_MULTIMEM handle of rank %d of this job", r, r);
        h.server[sizeof h.server - 1] = 0;
        c->peer_server[r] = h.server;
        if (r == rank) continue;
        int fd = -1;
        if (int rc = vmm::fetch_fd(c->peer_server[r], 'A', &fd)) return rc;
        const int rc = vmm::import_fd(c->device, fd, (size_t)h.bytes, &c->vm_peer[r]);
        close(fd);
        if (rc) return rc;
        c->base[r] = reinterpret_cast<float*>(c->vm_peer[r].va);
    }
    if (rank != 0) {
        int fd = -1;
        if (int rc = vmm::fetch_fd(c->peer_server[0], 'M', &fd)) return rc;
        const int rc = vmm::mc_import(fd, c->mc_bytes, &c->mc);
        close(fd);
        if (rc) return rc;
    }
    // every rank adds its device now; the memory is bound at the first collective call, i.e. after the
    // host barrier the caller places behind this function (binding needs every device to have been added)
    if (int rc = vmm::mc_add_device(&c->mc, c->device)) return rc;
    c->imported = true;
    return B200PS_OK;
}

extern "C" int b200ps_ctx_sync(b200ps_ctx_t* c, void* stream) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    B200PS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B200PS_CUDA(cudaStreamSynchronize(c->s_in));
    B200PS_CUDA(cudaStreamSynchronize(c->s_out));
    return check_alive(c);
}

extern "C" const b200ps_plan_t* b200ps_ctx_plan(const b200ps_ctx_t* c) { return c ? c->plan.get() : nullptr; }
extern "C" uint64_t b200ps_ctx_step_count(const b200ps_ctx_t* c) { return c ? c->t : 0; }

extern "C" uint64_t b200ps_ctx_algorithmic_bytes_per_step(const b200ps_ctx_t* c) {
    if (!c) return 0;
    const uint64_t k = 4ull * (uint64_t)c->plan->W + (c->o.optimizer == B200PS_OPT_SGD ? 8 : c->o.optimizer == B200PS_OPT_MOMENTUM ? 16 : 24);
    uint64_t b = 0;
    for (const auto& L : c->plan->launches) b += k * (L.end - L.begin);
    return b;
}

extern "C" int b200ps_ctx_set_tuning(b200ps_ctx_t* c, int threads, int ctas_per_sm, int unroll) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    B200PS_REQUIRE((!threads || (threads >= 32 && threads <= 1024 && threads % 32 == 0)) &&
                       (!ctas_per_sm || (ctas_per_sm >= 1 && ctas_per_sm <= 32)) && (!unroll || unroll == 1 || unroll == 2 || unroll == 4),
                   B200PS_ERR_RANGE, "launch geometry: threads a multiple of 32 in 32..1024, ctas_per_sm 1..32, unroll 1|2|4");
    if (threads) c->tune.threads = threads;
    if (ctas_per_sm) c->tune.ctas_per_sm = ctas_per_sm;
    if (unroll) c->tune.unroll = unroll;
    drop_graph(c);  // the captured grids are stale
    return B200PS_OK;
}

extern "C" int b200ps_ctx_set_graph(b200ps_ctx_t* c, int on) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    c->o.use_graph = on != 0;
    return B200PS_OK;
}

extern "C" int b200ps_ctx_set_profile(b200ps_ctx_t* c, int on) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    c->profile = on != 0;
    return B200PS_OK;
}

extern "C" int b200ps_ctx_phase_ms(b200ps_ctx_t* c, float* fold_ms, float* apply_ms, float* barrier_ms) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    B200PS_CUDA(cudaDeviceSynchronize());
    float* out[3] = {fold_ms, apply_ms, barrier_ms};
    // per step: fold 1 span, apply 1 span, barrier 2 spans -> report per-step sums
    const size_t steps = c->spans[1].size();
    for (int k = 0; k < 3; ++k) {
        double sum = 0;
        for (const auto& s : c->spans[k]) {
            float ms = 0;
            B200PS_CUDA(cudaEventElapsedTime(&ms, s.a, s.b));
            sum += ms;
        }
        if (out[k]) *out[k] = steps ? (float)(sum / (double)steps) : 0.f;
        c->spans[k].clear();
    }
    for (cudaEvent_t e : c->event_pool) cudaEventDestroy(e);
    c->event_pool.clear();
    return B200PS_OK;
}

extern "C" int b200ps_multicast_supported(void) { return vmm::available() ? 1 : 0; }

// ---- parameters ---------------------------------------------------------------------------------
extern "C" int b200ps_dense_init(b200ps_ctx_t* c, const float* h_params, void* stream) {
    if (int rc = check_alive(c)) return rc;
    B200PS_REQUIRE(h_params, B200PS_ERR_INVALID, "null parameters");
    if (int rc = wait_copy_out(c, (cudaStream_t)stream)) return rc;
    B200PS_CUDA(cudaMemcpyAsync(c->arena, h_params, (size_t)c->plan->P * 4, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return zero_state(c, (cudaStream_t)stream);
}

extern "C" int b200ps_dense_init_uniform(b200ps_ctx_t* c, uint64_t seed, void* stream) {
    if (int rc = check_alive(c)) return rc;
    if (int rc = b200ps_fill_uniform(c->arena, c->plan->P, seed, 0, 0, 0, stream)) return rc;
    return zero_state(c, (cudaStream_t)stream);
}

extern "C" int b200ps_dense_replica(b200ps_ctx_t* c, float** d_params) {
    B200PS_REQUIRE(c && d_params, B200PS_ERR_INVALID, "null argument");
    *d_params = c->arena;
    return B200PS_OK;
}

extern "C" int b200ps_dense_grad_buffer(b200ps_ctx_t* c, int worker, float** d_grad) {
    B200PS_REQUIRE(c && d_grad, B200PS_ERR_INVALID, "null argument");
    const auto& p = *c->plan;
    B200PS_REQUIRE(worker >= 0 && worker < p.W && b200ps_placement_gpu(worker, p.world) == p.rank, B200PS_ERR_RANGE,
                   "compute task %d is not placed on rank %d", worker, p.rank);
    *d_grad = c->arena + p.grad_offset(worker).second;
    return B200PS_OK;
}

extern "C" int b200ps_ctx_arena_base(b200ps_ctx_t* c, int of_rank, float** d_base) {
    B200PS_REQUIRE(c && d_base, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(of_rank >= 0 && of_rank < c->plan->world, B200PS_ERR_RANGE, "no rank %d", of_rank);
    B200PS_REQUIRE(c->base[of_rank], B200PS_ERR_STATE, "rank %d's arena is not mapped yet (b200ps_ctx_import_handles)", of_rank);
    *d_base = c->base[of_rank];
    return B200PS_OK;
}

extern "C" int b200ps_dense_state(b200ps_ctx_t* c, int shard, float** d_first, float** d_second, uint64_t* n) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    const auto& p = *c->plan;
    for (size_t i = 0; i < p.launches.size(); ++i)
        if (p.launches[i].shard == shard) {
            if (d_first) *d_first = c->s1[i];
            if (d_second) *d_second = c->s2[i];
            if (n) *n = p.launches[i].end - p.launches[i].begin;
            return B200PS_OK;
        }
    return fail(B200PS_ERR_RANGE, "shard %d is not placed on rank %d", shard, p.rank);
}

// ---- the step -----------------------------------------------------------------------------------
extern "C" int b200ps_dense_push(b200ps_ctx_t* c, int worker, const float* h_grad, void* stream) {
    if (int rc = check_alive(c)) return rc;
    float* d = nullptr;
    if (int rc = b200ps_dense_grad_buffer(c, worker, &d)) return rc;
    B200PS_REQUIRE(h_grad, B200PS_ERR_INVALID, "null gradient");
    B200PS_CUDA(cudaMemcpyAsync(d, h_grad, (size_t)c->plan->P * 4, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return B200PS_OK;
}

extern "C" int b200ps_dense_pull(b200ps_ctx_t* c, int worker, float* h_params, void* stream) {
    if (int rc = check_alive(c)) return rc;
    float* d = nullptr;
    if (int rc = b200ps_dense_grad_buffer(c, worker, &d)) return rc;  // validates that the worker is local
    B200PS_REQUIRE(h_params, B200PS_ERR_INVALID, "null destination");
    B200PS_CUDA(cudaMemcpyAsync(h_params, c->arena, (size_t)c->plan->P * 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return B200PS_OK;
}

extern "C" int b200ps_dense_step(b200ps_ctx_t* c, void* stream) {
    if (int rc = check_alive(c)) return rc;
    if (int rc = need_peers(c)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = wait_copy_out(c, st)) return rc;
    if (c->o.use_graph && !c->profile) {
        if (!c->graph) {
            // recorded on a stream of the context (the caller's may be the legacy default stream, which
            // cannot capture); recording executes nothing, and the graph is launched on the caller's stream
            cudaGraph_t g = nullptr;
            B200PS_CUDA(cudaStreamBeginCapture(c->s_in, cudaStreamCaptureModeThreadLocal));
            const int rc = enqueue_step(c, c->s_in);
            const cudaError_t e = cudaStreamEndCapture(c->s_in, &g);
            if (rc) {
                if (g) cudaGraphDestroy(g);
                return rc;
            }
            B200PS_CUDA(e);
            {
                size_t n = 0;
                cudaGraphGetNodes(g, nullptr, &n);
                std::vector<cudaGraphNode_t> nodes(n);
                if (n) cudaGraphGetNodes(g, nodes.data(), &n);
                c->graph_kernels = 0;
                for (cudaGraphNode_t nd : nodes) {
                    cudaGraphNodeType ty;
                    if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel) c->graph_kernels += 1;
                    // the nodes whose step size changes every step
                    if (c->adam() && dense_node_opt(nd) == OPT_ADAM) c->graph_adam_nodes.push_back(nd);
                }
            }
            const cudaError_t e2 = cudaGraphInstantiate(&c->graph, g, 0);
            // the graph object must outlive the exec for node handles to stay meaningful: keep it until drop
            c->graph_src.reset(g);
            B200PS_CUDA(e2);
        } else if (c->adam()) {
            const float lr_t = lr_for_step(c, c->t + 1);
            for (cudaGraphNode_t nd : c->graph_adam_nodes) B200PS_CUDA(dense_node_set_lr(c->graph, nd, lr_t));
        }
        B200PS_CUDA(cudaGraphLaunch(c->graph, st));
        g_launches.fetch_add(c->graph_kernels, std::memory_order_relaxed);  // the kernels one replay launches
    } else {
        if (int rc = enqueue_step(c, st)) return rc;
    }
    c->t += 1;
    return B200PS_OK;
}

extern "C" int b200ps_dense_step_async_ordered(b200ps_ctx_t* c, void* stream) {
    if (int rc = check_alive(c)) return rc;
    if (int rc = need_peers(c)) return rc;
    const auto& p = *c->plan;
    B200PS_REQUIRE(!p.two_level(), B200PS_ERR_STATE, "the ordered-async step needs B200PS_AGG_FLAT (every launch folds one worker)");
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = wait_copy_out(c, st)) return rc;
    if (int rc = barrier(c, st)) return rc;  // every worker's gradient is in place
    for (int w = 0; w < p.W; ++w) {
        for (size_t i = 0; i < p.launches.size(); ++i)
            if (int rc = launch_apply(c, i, p.launches[i].begin, p.launches[i].end, c->t + 1, st, w, w == p.W - 1, 1.0f)) return rc;
        c->t += 1;
    }
    return barrier(c, st);
}

namespace {

int ensure_chunk_events(b200ps_ctx* c, size_t n) {
    while (c->ev_in.size() < n) {
        cudaEvent_t a, b, d;
        B200PS_CUDA(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        B200PS_CUDA(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        B200PS_CUDA(cudaEventCreateWithFlags(&d, cudaEventDisableTiming));
        c->ev_in.push_back(a);
        c->ev_k.push_back(b);
        c->ev_out.push_back(d);
        c->ev_out_valid.push_back(0);
    }
    return B200PS_OK;
}

}  // namespace

extern "C" int b200ps_dense_step_host(b200ps_ctx_t* c, const float* const* h_grads, float* h_out, void* stream) {
    if (int rc = check_alive(c)) return rc;
    if (int rc = need_peers(c)) return rc;
    const auto& p = *c->plan;
    B200PS_REQUIRE(h_grads || p.local_workers.empty(), B200PS_ERR_INVALID, "null gradient list");
    for (size_t i = 0; i < p.local_workers.size(); ++i)
        B200PS_REQUIRE(h_grads[i], B200PS_ERR_INVALID, "host gradient of local compute task %zu is null", i);
    cudaStream_t st = (cudaStream_t)stream;
    const uint64_t chunk = c->o.host_chunk_elems;
    // this step's H2D may start once the previous step's kernels are done with the gradient buffers;
    // it does NOT wait for the previous step's D2H (the other PCIe direction)
    B200PS_CUDA(cudaEventRecord(c->ev_fork, st));
    B200PS_CUDA(cudaStreamWaitEvent(c->s_in, c->ev_fork, 0));
    if (p.world == 1) {
        // one GPU: H2D, kernel and D2H pipelined chunk by chunk over three streams
        size_t ci = 0;
        for (size_t i = 0; i < p.launches.size(); ++i) {
            const auto& L = p.launches[i];
            for (uint64_t lo = L.begin; lo < L.end; lo += chunk, ++ci) {
                const uint64_t hi = lo + chunk < L.end ? lo + chunk : L.end;
                if (int rc = ensure_chunk_events(c, ci + 1)) return rc;
                for (size_t k = 0; k < p.local_workers.size(); ++k)
                    B200PS_CUDA(cudaMemcpyAsync(c->arena + p.grad_offset(p.local_workers[k]).second + lo, h_grads[k] + lo,
                                                (size_t)(hi - lo) * 4, cudaMemcpyHostToDevice, c->s_in));
                B200PS_CUDA(cudaEventRecord(c->ev_in[ci], c->s_in));
                B200PS_CUDA(cudaStreamWaitEvent(st, c->ev_in[ci], 0));
                if (c->ev_out_valid[ci])  // the previous step's copy-out of this chunk has left the replica
                    B200PS_CUDA(cudaStreamWaitEvent(st, c->ev_out[ci], 0));
                if (int rc = launch_apply(c, i, lo, hi, c->t + 1, st)) return rc;
                if (h_out) {
                    B200PS_CUDA(cudaEventRecord(c->ev_k[ci], st));
                    B200PS_CUDA(cudaStreamWaitEvent(c->s_out, c->ev_k[ci], 0));
                    B200PS_CUDA(cudaMemcpyAsync(h_out + lo, c->arena + lo, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, c->s_out));
                    B200PS_CUDA(cudaEventRecord(c->ev_out[ci], c->s_out));
                    c->ev_out_valid[ci] = 1;
                    c->out_busy = true;
                }
            }
        }
        c->t += 1;
        return B200PS_OK;
    }
    // several GPUs: copy-in pipelined with the local fold chunk by chunk; then barrier, fused
    // applies, barrier; the copy-out of the whole replica overlaps the next step's copy-in.
    size_t ci = 0;
    for (uint64_t lo = 0; lo < p.P; lo += chunk, ++ci) {
        const uint64_t hi = lo + chunk < p.P ? lo + chunk : p.P;
        if (int rc = ensure_chunk_events(c, ci + 1)) return rc;
        for (size_t k = 0; k < p.local_workers.size(); ++k)
            B200PS_CUDA(cudaMemcpyAsync(c->arena + p.grad_offset(p.local_workers[k]).second + lo, h_grads[k] + lo,
                                        (size_t)(hi - lo) * 4, cudaMemcpyHostToDevice, c->s_in));
        B200PS_CUDA(cudaEventRecord(c->ev_in[ci], c->s_in));
        B200PS_CUDA(cudaStreamWaitEvent(st, c->ev_in[ci], 0));
        if (int rc = launch_fold(c, lo, hi, st)) return rc;
    }
    // peers' applies of THIS step write into my replica: my previous copy-out must have left it
    // before anyone passes the first barrier
    if (int rc = wait_copy_out(c, st)) return rc;
    if (int rc = barrier(c, st)) return rc;
    for (size_t i = 0; i < p.launches.size(); ++i)
        if (int rc = launch_apply(c, i, p.launches[i].begin, p.launches[i].end, c->t + 1, st)) return rc;
    if (int rc = barrier(c, st)) return rc;
    if (h_out) {
        B200PS_CUDA(cudaEventRecord(c->ev_join, st));
        B200PS_CUDA(cudaStreamWaitEvent(c->s_out, c->ev_join, 0));
        B200PS_CUDA(cudaMemcpyAsync(h_out, c->arena, (size_t)p.P * 4, cudaMemcpyDeviceToHost, c->s_out));
        c->out_busy = true;
    }
    c->t += 1;
    return B200PS_OK;
}

extern "C" int b200ps_dense_host_sync(b200ps_ctx_t* c) {
    B200PS_REQUIRE(c, B200PS_ERR_INVALID, "null context");
    B200PS_CUDA(cudaStreamSynchronize(c->s_out));
    return check_alive(c);
}

// ---- inspection, checkpoints ----------------------------------------------------------------------
extern "C" int b200ps_dense_dump(b200ps_ctx_t* c, float* h_params, float* h_first, float* h_second, void* stream) {
    if (int rc = b200ps_ctx_sync(c, stream)) return rc;
    const auto& p = *c->plan;
    if (h_params) B200PS_CUDA(cudaMemcpy(h_params, c->arena, (size_t)p.P * 4, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < p.launches.size(); ++i) {
        const auto& L = p.launches[i];
        const size_t bytes = (size_t)(L.end - L.begin) * 4;
        if (h_first && c->s1[i]) B200PS_CUDA(cudaMemcpy(h_first + L.begin, c->s1[i], bytes, cudaMemcpyDeviceToHost));
        if (h_second && c->s2[i]) B200PS_CUDA(cudaMemcpy(h_second + L.begin, c->s2[i], bytes, cudaMemcpyDeviceToHost));
    }
    return B200PS_OK;
}

extern "C" int b200ps_dense_save(b200ps_ctx_t* c, const char* directory, void* stream) {
    B200PS_REQUIRE(directory, B200PS_ERR_INVALID, "null directory");
    if (int rc = b200ps_ctx_sync(c, stream)) return rc;
    const auto& p = *c->plan;
    if (mkdir(directory, 0777) != 0 && errno != EEXIST) return fail(B200PS_ERR_INVALID, "cannot create %s: %s", directory, strerror(errno));
    const std::string path = ckpt_path(c, directory, p.rank), tmp = path + ".tmp" + std::to_string((long)getpid());
    FILE* f = fopen(tmp.c_str(), "wb");
    B200PS_REQUIRE(f, B200PS_ERR_INVALID, "cannot write %s: %s", tmp.c_str(), strerror(errno));
    CkptHeader h{};
    memcpy(h.magic, "B200PSD2", 8);
    h.P = p.P;
    h.t = c->t;
    h.S = p.S;
    h.W = p.W;
    h.optimizer = c->o.optimizer;
    h.rank = p.rank;
    h.world = p.world;
    h.n_shards = (int32_t)p.launches.size();
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    std::vector<float> buf;
    for (size_t i = 0; ok && i < p.launches.size(); ++i) {
        const auto& L = p.launches[i];
        const size_t n = (size_t)(L.end - L.begin);
        CkptShard sh{L.shard, (c->s1[i] ? 1 : 0) + (c->s2[i] ? 1 : 0), L.begin, L.end};
        ok = fwrite(&sh, sizeof sh, 1, f) == 1;
        buf.resize(n);
        const float* srcs[3] = {c->arena + L.begin, c->s1[i], c->s2[i]};
        for (int k = 0; ok && k < 3; ++k) {
            if (!srcs[k] || !n) continue;
            if (cudaMemcpy(buf.data(), srcs[k], n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) ok = false;
            ok = ok && fwrite(buf.data(), 4, n, f) == n;
        }
    }
    ok = (fflush(f) == 0) && ok;
    ok = (fsync(fileno(f)) == 0) && ok;
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path.c_str()) != 0) {
        unlink(tmp.c_str());
        return fail(B200PS_ERR_INVALID, "writing %s failed: %s", path.c_str(), strerror(errno));
    }
    return B200PS_OK;
}

extern "C" int b200ps_dense_load(b200ps_ctx_t* c, const char* directory, void* stream) {
    B200PS_REQUIRE(directory, B200PS_ERR_INVALID, "null directory");
    if (int rc = b200ps_ctx_sync(c, stream)) return rc;
    const auto& p = *c->plan;
    std::vector<float> buf;
    for (int r = 0; r < p.world; ++r) {
        const std::string path = ckpt_path(c, directory, r);
        FILE* f = fopen(path.c_str(), "rb");
        B200PS_REQUIRE(f, B200PS_ERR_INVALID, "cannot read %s: %s", path.c_str(), strerror(errno));
        std::unique_ptr<FILE, int (*)(FILE*)> guard(f, fclose);
        CkptHeader h{};
        B200PS_REQUIRE(fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, "B200PSD2", 8) == 0, B200PS_ERR_INVALID,
                       "%s is not a b200ps dense checkpoint (format B200PSD2)", path.c_str());
        B200PS_REQUIRE(h.P == p.P && h.S == p.S && h.W == p.W && h.optimizer == c->o.optimizer && h.world == p.world && h.rank == r,
                       B200PS_ERR_INVALID,
                       "%s was written for a different job layout (P_pad %llu, %d shards, %d workers, %s, rank %d of %d)", path.c_str(),
                       (unsigned long long)h.P, h.S, h.W, opt_name(h.optimizer), h.rank, h.world);
        for (int k = 0; k < h.n_shards; ++k) {
            CkptShard sh{};
            B200PS_REQUIRE(fread(&sh, sizeof sh, 1, f) == 1, B200PS_ERR_INVALID, "%s is truncated", path.c_str());
            B200PS_REQUIRE(sh.shard >= 0 && sh.shard < p.S && sh.begin == p.slices[(size_t)sh.shard].first &&
                               sh.end == p.slices[(size_t)sh.shard].second,
                           B200PS_ERR_INVALID, "%s: shard %d does not match this job's shard map", path.c_str(), sh.shard);
            const size_t n = (size_t)(sh.end - sh.begin);
            buf.resize(n);
            float* dst[3] = {c->arena + sh.begin, nullptr, nullptr};
            if (r == p.rank)
                for (size_t i = 0; i < p.launches.size(); ++i)
                    if (p.launches[i].shard == sh.shard) dst[1] = c->s1[i], dst[2] = c->s2[i];
            for (int a = 0; a < 1 + sh.n_state; ++a) {
                if (!n) continue;
                B200PS_REQUIRE(fread(buf.data(), 4, n, f) == n, B200PS_ERR_INVALID, "%s is truncated", path.c_str());
                if (dst[a]) B200PS_CUDA(cudaMemcpy(dst[a], buf.data(), n * 4, cudaMemcpyHostToDevice));
            }
        }
        if (r == p.rank) c->t = h.t;
    }
    return B200PS_OK;
}
