// table.cu -- one embedding table of the job as an object the library owns (include/b200ps.h "Table
// group"; SURVEY.md 8(b) sparse group, rows a10..a12, 8(e)).
//
// Reference: none -- TensorFlow's partitioned embedding variable on the PS tasks, embedding_lookup and
// the sparse optimizer ops (k8s-operator.md:6 gives only the roles).  The rows are partitioned by TF
// "div" over the S PS shards; shard s lives on GPU s mod G, compute task w on GPU w mod G.  Each rank
// owns ONE arena, exported to the other ranks (CUDA IPC):
//
//   [ rows of local shard 0 | local shard 1 | ... | (Adagrad accumulators, same order) |
//     per local compute task: ids int64[B] | gradient rows fp32[B, dim] | pulled rows fp32[B, dim] |
//     per REMOTE compute task: staging fp32[B, dim] for the gradient rows it routes here | barrier words ]
//
// Every rank computes every rank's layout, so after the one-time handle exchange the kernels get
// plain base pointers and the routing all-to-alls of a classic PS are fused into them (csrc/sparse.cu):
// ids are streamed from wherever the workers wrote them, rows cross NVLink only as posted stores.
//   pull  barrier -> serve (every owner serves its rows into the requesting tasks' buffers) -> barrier
//   push  route (rows owned elsewhere go to the owners' staging slots) -> barrier ->
//         sort_segment -> apply -> barrier
//   step  (pull + push of the same ids as one unit) the push plan needs the ids only, so sort_segment
//         runs on a second, high-priority stream BESIDE serve and route and joins before apply:
//         barrier -> { serve -> route -> barrier | sort_segment } -> apply -> barrier
#include <cuda_runtime.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "common.h"
#include "cuda_util.h"
#include "peer_sync.cuh"

namespace b200ps {
namespace {

constexpr uint64_t NONE = ~0ull;
uint64_t up256(uint64_t x) { return (x + 255) / 256 * 256; }

struct Layout {
    uint64_t rows = 0, B = 0;
    int dim = 0, S = 0, W = 0, world = 1;
    bool adagrad = false;
    std::vector<std::pair<uint64_t, uint64_t>> shard_range;  // rows of shard s
    std::vector<uint64_t> shard_off, accum_off;              // per shard, byte offset in its owner's arena
    std::vector<uint64_t> ids_off, grad_off, out_off;        // per compute task, in its GPU's arena
    std::vector<std::vector<uint64_t>> stage_off;            // [rank][task]: staging slot on `rank` for a REMOTE task
    std::vector<uint64_t> flag_off, arena_bytes;             // per rank
    int shard_rank(int s) const { return s % world; }
    int worker_rank(int w) const { return w % world; }
};

int build_layout(uint64_t rows, int dim, int S, int W, uint64_t B, int world, bool adagrad, Layout* L) {
    B200PS_REQUIRE(rows >= 1 && rows < (1ull << 32), B200PS_ERR_RANGE, "rows must be in 1 .. 2^32 - 1");
    B200PS_REQUIRE(dim >= 4 && dim <= 4096 && dim % 4 == 0, B200PS_ERR_RANGE, "dim must be a multiple of 4 in 4..4096");
    B200PS_REQUIRE(S >= 1 && S <= B200PS_MAX_WORKERS && W >= 1 && W <= B200PS_MAX_WORKERS, B200PS_ERR_RANGE,
                   "1..%d shards and compute tasks", B200PS_MAX_WORKERS);
    B200PS_REQUIRE(world >= 1 && world <= SYNC_MAX_RANKS, B200PS_ERR_RANGE, "1..%d ranks", SYNC_MAX_RANKS);
    B200PS_REQUIRE(B >= 1 && B * (uint64_t)W < (1ull << 32), B200PS_ERR_RANGE, "ids_per_worker * workers must be below 2^32");
    L->rows = rows;
    L->dim = dim;
    L->S = S;
    L->W = W;
    L->B = B;
    L->world = world;
    L->adagrad = adagrad;
    L->shard_range.resize((size_t)S);
    for (int s = 0; s < S; ++s) b200ps_row_range(rows, S, s, &L->shard_range[(size_t)s].first, &L->shard_range[(size_t)s].second);
    L->shard_off.assign((size_t)S, NONE);
    L->accum_off.assign((size_t)S, NONE);
    L->ids_off.assign((size_t)W, NONE);
    L->grad_off.assign((size_t)W, NONE);
    L->out_off.assign((size_t)W, NONE);
    L->stage_off.assign((size_t)world, std::vector<uint64_t>((size_t)W, NONE));
    L->flag_off.assign((size_t)world, 0);
    L->arena_bytes.assign((size_t)world, 0);
    const uint64_t row_bytes = (uint64_t)dim * 4;
    for (int r = 0; r < world; ++r) {
        uint64_t off = 0;
        for (int s = 0; s < S; ++s)
            if (L->shard_rank(s) == r) {
                L->shard_off[(size_t)s] = off;
                off += up256((L->shard_range[(size_t)s].second - L->shard_range[(size_t)s].first) * row_bytes);
            }
        if (adagrad)
            for (int s = 0; s < S; ++s)
                if (L->shard_rank(s) == r) {
                    L->accum_off[(size_t)s] = off;
                    off += up256((L->shard_range[(size_t)s].second - L->shard_range[(size_t)s].first) * row_bytes);
                }
        for (int w = 0; w < W; ++w)
            if (L->worker_rank(w) == r) {
                L->ids_off[(size_t)w] = off;
                off += up256(B * 8);
                L->grad_off[(size_t)w] = off;
                off += up256(B * row_bytes);
                L->out_off[(size_t)w] = off;
                off += up256(B * row_bytes);
            }
        for (int w = 0; w < W; ++w)
            if (L->worker_rank(w) != r) {
                L->stage_off[(size_t)r][(size_t)w] = off;  // rows task w routes to GPU r
                off += up256(B * row_bytes);
            }
        L->flag_off[(size_t)r] = off;  // barrier words behind everything else
        off += up256(B200PS_BARRIER_WORDS * 4);
        L->arena_bytes[(size_t)r] = off;
    }
    return B200PS_OK;
}

struct TableState {
    uint32_t epoch, abort, pad[2];
};

struct TblHeader {
    char magic[8];  // "B200PST2"
    uint64_t rows;
    int32_t dim, S, optimizer, rank, world, n_shards;
};

}  // namespace
}  // namespace b200ps

using namespace b200ps;

struct b200ps_table {
    Layout L;
    b200ps_table_options_t o{};
    b200ps_table_geom_t geom{};
    int rank = 0, world = 1, device = 0;
    char* arena = nullptr;
    char* base[SYNC_MAX_RANKS] = {nullptr};
    bool imported = false;
    void* workspace = nullptr;
    size_t workspace_bytes = 0;
    TableState* st = nullptr;
    uint32_t* host_abort = nullptr;
    uint32_t* host_abort_dev = nullptr;
    std::vector<int> local_shards, local_workers;
    // pointer tables the kernels take (host arrays of device pointers)
    std::vector<float*> shard_ptrs, accum_ptrs, out_ptrs;
    std::vector<const float*> shard_cptrs, grad_ptrs;
    std::vector<const int64_t*> ids_ptrs;
    std::vector<std::vector<float*>> route_dst;  // per local task: staging slot on every GPU (null for this one)
    cudaStream_t s_cap = nullptr;
    cudaStream_t s_side = nullptr;                 // the plan of a step is made here, beside the pull
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaGraphExec_t graph = nullptr;
    uint64_t graph_kernels = 0;
    bool profile = false;
    struct Span { cudaEvent_t a, b; };
    std::vector<Span> spans[3];  // serve, sort_segment (incl. route + barrier), apply
    std::vector<cudaEvent_t> event_pool;
};

namespace {

int alive(const b200ps_table* t) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    B200PS_REQUIRE(*(volatile uint32_t*)t->host_abort == 0, B200PS_ERR_STATE,
                   "a peer barrier of this table timed out (some rank never arrived)");
    B200PS_REQUIRE(t->world == 1 || t->imported, B200PS_ERR_STATE, "world = %d: call b200ps_table_import_handles first", t->world);
    return B200PS_OK;
}

void build_pointers(b200ps_table* t) {
    const Layout& L = t->L;
    t->shard_ptrs.assign((size_t)L.S, nullptr);
    t->accum_ptrs.assign((size_t)L.S, nullptr);
    t->shard_cptrs.assign((size_t)L.S, nullptr);
    for (int s = 0; s < L.S; ++s) {
        char* b = t->base[L.shard_rank(s)];
        if (!b) continue;
        t->shard_ptrs[(size_t)s] = reinterpret_cast<float*>(b + L.shard_off[(size_t)s]);
        t->shard_cptrs[(size_t)s] = t->shard_ptrs[(size_t)s];
        if (L.adagrad) t->accum_ptrs[(size_t)s] = reinterpret_cast<float*>(b + L.accum_off[(size_t)s]);
    }
    t->ids_ptrs.assign((size_t)L.W, nullptr);
    t->grad_ptrs.assign((size_t)L.W, nullptr);
    t->out_ptrs.assign((size_t)L.W, nullptr);
    for (int w = 0; w < L.W; ++w) {
        char* b = t->base[L.worker_rank(w)];
        if (b) {
            t->ids_ptrs[(size_t)w] = reinterpret_cast<const int64_t*>(b + L.ids_off[(size_t)w]);
            t->out_ptrs[(size_t)w] = reinterpret_cast<float*>(b + L.out_off[(size_t)w]);
        }
        // apply reads local memory only: a local task's own buffer, or this GPU's staging slot for a remote one
        const uint64_t off = L.worker_rank(w) == t->rank ? L.grad_off[(size_t)w] : L.stage_off[(size_t)t->rank][(size_t)w];
        t->grad_ptrs[(size_t)w] = reinterpret_cast<const float*>(t->arena + off);
    }
    t->route_dst.clear();
    for (int w : t->local_workers) {
        std::vector<float*> dst((size_t)t->world, nullptr);
        for (int o = 0; o < t->world; ++o)
            if (o != t->rank && t->base[o]) dst[(size_t)o] = reinterpret_cast<float*>(t->base[o] + L.stage_off[(size_t)o][(size_t)w]);
        t->route_dst.push_back(std::move(dst));
    }
}

int tbarrier(b200ps_table* t, cudaStream_t st) {
    if (t->world == 1) return B200PS_OK;
    PeerBarrierArgs a{};
    for (int r = 0; r < t->world; ++r) a.flags[r] = reinterpret_cast<uint32_t*>(t->base[r] + t->L.flag_off[(size_t)r]);
    a.epoch = &t->st->epoch;
    a.abort = &t->st->abort;
    a.host_abort = t->host_abort_dev;
    a.rank = t->rank;
    a.world = t->world;
    peer_barrier_epoch_kernel<<<1, 32, 0, st>>>(a);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

struct TMark {
    b200ps_table* t;
    int phase;
    cudaStream_t st;
    cudaEvent_t a = nullptr;
    TMark(b200ps_table* t_, int phase_, cudaStream_t st_) : t(t_), phase(phase_), st(st_) {
        if (t->profile && cudaEventCreate(&a) == cudaSuccess) {
            t->event_pool.push_back(a);
            cudaEventRecord(a, st);
        }
    }
    ~TMark() {
        cudaEvent_t b = nullptr;
        if (a && cudaEventCreate(&b) == cudaSuccess) {
            t->event_pool.push_back(b);
            cudaEventRecord(b, st);
            t->spans[phase].push_back({a, b});
        }
    }
};

int do_serve(b200ps_table* t, cudaStream_t st) {
    return b200ps_sparse_serve(t->out_ptrs.data(), t->shard_cptrs.data(), t->ids_ptrs.data(), t->L.W, t->L.B, &t->geom, st);
}

int do_route(b200ps_table* t, cudaStream_t st) {
    if (t->world == 1) return B200PS_OK;
    for (size_t i = 0; i < t->local_workers.size(); ++i) {
        const int w = t->local_workers[i];
        if (int rc = b200ps_sparse_route(t->grad_ptrs[(size_t)w], t->ids_ptrs[(size_t)w], t->L.B, &t->geom, t->route_dst[i].data(), st)) return rc;
    }
    return B200PS_OK;
}

int do_sort(b200ps_table* t, cudaStream_t st) {
    return b200ps_sparse_sort_segment(t->ids_ptrs.data(), t->L.W, t->L.B, &t->geom, t->workspace, t->workspace_bytes, st);
}

int do_apply(b200ps_table* t, cudaStream_t st) {
    if (t->o.optimizer == B200PS_OPT_ADAGRAD)
        return b200ps_sparse_apply_adagrad(t->shard_ptrs.data(), t->accum_ptrs.data(), t->grad_ptrs.data(), t->L.W, t->L.B, &t->geom,
                                           (float)t->o.lr, t->workspace, st);
    return b200ps_sparse_apply_sgd(t->shard_ptrs.data(), t->grad_ptrs.data(), t->L.W, t->L.B, &t->geom, (float)t->o.lr, t->workspace, st);
}

int enqueue_pull(b200ps_table* t, cudaStream_t st, bool closing_barrier) {
    if (int rc = tbarrier(t, st)) return rc;  // every task's ids are written
    {
        TMark m(t, 0, st);
        if (int rc = do_serve(t, st)) return rc;
    }
    return closing_barrier ? tbarrier(t, st) : B200PS_OK;  // every owner has delivered its rows
}

int enqueue_push(b200ps_table* t, cudaStream_t st) {
    {
        TMark m(t, 1, st);
        if (int rc = do_route(t, st)) return rc;
        if (int rc = tbarrier(t, st)) return rc;  // all ids written, every routed row has landed (and every served row too)
        if (int rc = do_sort(t, st)) return rc;
    }
    {
        TMark m(t, 2, st);
        if (int rc = do_apply(t, st)) return rc;
    }
    return tbarrier(t, st);  // every shard is updated before anyone pulls again
}

// pull + push of the same ids as one unit.  The barrier that closes the pull and the one that guards
// the push are the same one (the routed rows and the served rows land behind it together), and the
// push plan is made on the side stream while the rows are being served.  With phase timing on, the
// phases run one after the other so that each span means what its name says.
int enqueue_step(b200ps_table* t, cudaStream_t st) {
    if (t->profile) {
        if (int rc = enqueue_pull(t, st, false)) return rc;
        return enqueue_push(t, st);
    }
    if (int rc = tbarrier(t, st)) return rc;  // every task's ids are written
    B200PS_CUDA(cudaEventRecord(t->ev_fork, st));
    B200PS_CUDA(cudaStreamWaitEvent(t->s_side, t->ev_fork, 0));
    int rc = do_sort(t, t->s_side);
    const cudaError_t e1 = cudaEventRecord(t->ev_join, t->s_side);
    if (!rc) rc = do_serve(t, st);
    if (!rc) rc = do_route(t, st);
    if (!rc) rc = tbarrier(t, st);  // every served and every routed row has landed
    const cudaError_t e2 = cudaStreamWaitEvent(st, t->ev_join, 0);  // always joined: a capture must not end with a loose branch
    if (rc) return rc;
    B200PS_CUDA(e1);
    B200PS_CUDA(e2);
    if (int rc2 = do_apply(t, st)) return rc2;
    return tbarrier(t, st);  // every shard is updated before anyone pulls again
}

std::string tbl_path(const b200ps_table* t, const char* dir) {
    return std::string(dir) + "/table-rank" + std::to_string(t->rank) + "-of-" + std::to_string(t->world) + ".b200ps";
}

}  // namespace

extern "C" int b200ps_table_options_default(b200ps_table_options_t* o) {
    B200PS_REQUIRE(o, B200PS_ERR_INVALID, "null argument");
    memset(o, 0, sizeof *o);
    o->optimizer = B200PS_OPT_SGD;
    o->lr = 0.05;
    o->accum_init = 0.1;
    return B200PS_OK;
}

extern "C" int b200ps_table_layout(uint64_t rows, int dim, int num_shards, int num_workers, uint64_t ids_per_worker, int world,
                                   int adagrad, int of_rank, uint64_t* arena_bytes, uint64_t* flag_offset, uint64_t* shard_offset,
                                   uint64_t* accum_offset, uint64_t* ids_offset, uint64_t* grad_offset, uint64_t* pulled_offset,
                                   uint64_t* staging_offset) {
    Layout L;
    if (int rc = build_layout(rows, dim, num_shards, num_workers, ids_per_worker, world, adagrad != 0, &L)) return rc;
    B200PS_REQUIRE(of_rank >= 0 && of_rank < world, B200PS_ERR_RANGE, "no rank %d", of_rank);
    if (arena_bytes) *arena_bytes = L.arena_bytes[(size_t)of_rank];
    if (flag_offset) *flag_offset = L.flag_off[(size_t)of_rank];
    for (int s = 0; s < num_shards; ++s) {
        const bool here = L.shard_rank(s) == of_rank;
        if (shard_offset) shard_offset[s] = here ? L.shard_off[(size_t)s] : NONE;
        if (accum_offset) accum_offset[s] = here ? L.accum_off[(size_t)s] : NONE;
    }
    for (int w = 0; w < num_workers; ++w) {
        const bool here = L.worker_rank(w) == of_rank;
        if (ids_offset) ids_offset[w] = here ? L.ids_off[(size_t)w] : NONE;
        if (grad_offset) grad_offset[w] = here ? L.grad_off[(size_t)w] : NONE;
        if (pulled_offset) pulled_offset[w] = here ? L.out_off[(size_t)w] : NONE;
        if (staging_offset) staging_offset[w] = L.stage_off[(size_t)of_rank][(size_t)w];
    }
    return B200PS_OK;
}

extern "C" int b200ps_table_create(const b200ps_cluster_t* cluster, uint64_t rows, int dim, uint64_t ids_per_worker, int rank, int world,
                                   const b200ps_table_options_t* opt, b200ps_table_t** out) {
    B200PS_REQUIRE(cluster && out, B200PS_ERR_INVALID, "null argument");
    b200ps_table_options_t o;
    b200ps_table_options_default(&o);
    if (opt) o = *opt;
    B200PS_REQUIRE(o.optimizer == B200PS_OPT_SGD || o.optimizer == B200PS_OPT_ADAGRAD, B200PS_ERR_INVALID,
                   "a table is updated with B200PS_OPT_SGD or B200PS_OPT_ADAGRAD");
    B200PS_REQUIRE(o.lr > 0, B200PS_ERR_RANGE, "lr must be positive");
    B200PS_REQUIRE(rank >= 0 && rank < world, B200PS_ERR_RANGE, "rank %d is outside 0..%d", rank, world - 1);
    auto t = std::make_unique<b200ps_table>();
    const int S = b200ps_cluster_num_shards(cluster), W = b200ps_cluster_num_workers(cluster);
    if (int rc = build_layout(rows, dim, S, W, ids_per_worker, world, o.optimizer == B200PS_OPT_ADAGRAD, &t->L)) return rc;
    t->o = o;
    t->rank = rank;
    t->world = world;
    t->geom.rows = rows;
    t->geom.dim = dim;
    t->geom.nshards = S;
    t->geom.rank = rank;
    t->geom.world = world;
    B200PS_CUDA(cudaGetDevice(&t->device));
    for (int s = 0; s < S; ++s)
        if (t->L.shard_rank(s) == rank) t->local_shards.push_back(s);
    for (int w = 0; w < W; ++w)
        if (t->L.worker_rank(w) == rank) t->local_workers.push_back(w);
    B200PS_CUDA(cudaMalloc(&t->arena, t->L.arena_bytes[(size_t)rank]));
    t->base[rank] = t->arena;
    B200PS_CUDA(cudaMemset(t->arena + t->L.flag_off[(size_t)rank], 0, B200PS_BARRIER_WORDS * 4));
    t->workspace_bytes = b200ps_sparse_workspace_bytes((uint64_t)W * ids_per_worker, dim);
    B200PS_CUDA(cudaMalloc(&t->workspace, t->workspace_bytes));
    B200PS_CUDA(cudaMalloc(&t->st, sizeof(TableState)));
    B200PS_CUDA(cudaMemset(t->st, 0, sizeof(TableState)));
    B200PS_CUDA(cudaHostAlloc(&t->host_abort, sizeof(uint32_t), cudaHostAllocMapped));
    *t->host_abort = 0;
    B200PS_CUDA(cudaHostGetDevicePointer(&t->host_abort_dev, t->host_abort, 0));
    B200PS_CUDA(cudaStreamCreateWithFlags(&t->s_cap, cudaStreamNonBlocking));
    {
        int least = 0, greatest = 0;
        B200PS_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        // the sort is a chain of short, latency-bound kernels: its CTAs go first whenever a slot frees up
        B200PS_CUDA(cudaStreamCreateWithPriority(&t->s_side, cudaStreamNonBlocking, greatest));
        B200PS_CUDA(cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming));
        B200PS_CUDA(cudaEventCreateWithFlags(&t->ev_join, cudaEventDisableTiming));
    }
    if (t->L.adagrad) {  // TF's initial_accumulator_value
        for (int s : t->local_shards) {
            const size_t n = (size_t)(t->L.shard_range[(size_t)s].second - t->L.shard_range[(size_t)s].first) * (size_t)dim;
            std::vector<float> init(n, (float)o.accum_init);
            B200PS_CUDA(cudaMemcpy(t->arena + t->L.accum_off[(size_t)s], init.data(), n * 4, cudaMemcpyHostToDevice));
        }
    }
    build_pointers(t.get());
    B200PS_CUDA(cudaDeviceSynchronize());
    *out = t.release();
    return B200PS_OK;
}

extern "C" int b200ps_table_destroy(b200ps_table_t* t) {
    if (!t) return B200PS_OK;
    cudaDeviceSynchronize();
    if (t->graph) cudaGraphExecDestroy(t->graph);
    for (int r = 0; r < SYNC_MAX_RANKS; ++r)
        if (t->base[r] && t->base[r] != t->arena) cudaIpcCloseMemHandle(t->base[r]);
    for (cudaEvent_t e : t->event_pool) cudaEventDestroy(e);
    if (t->s_cap) cudaStreamDestroy(t->s_cap);
    if (t->s_side) cudaStreamDestroy(t->s_side);
    if (t->ev_fork) cudaEventDestroy(t->ev_fork);
    if (t->ev_join) cudaEventDestroy(t->ev_join);
    cudaFree(t->workspace);
    cudaFree(t->st);
    cudaFreeHost(t->host_abort);
    cudaFree(t->arena);
    delete t;
    return B200PS_OK;
}

extern "C" int b200ps_table_export_handle(b200ps_table_t* t, void* handle_out) {
    B200PS_REQUIRE(t && handle_out, B200PS_ERR_INVALID, "null argument");
    return b200ps_ipc_export(t->arena, handle_out);
}

extern "C" int b200ps_table_import_handles(b200ps_table_t* t, const void* handles) {
    B200PS_REQUIRE(t && (handles || t->world == 1), B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(!t->imported, B200PS_ERR_STATE, "handles were already imported");
    for (int r = 0; r < t->world; ++r) {
        if (r == t->rank) continue;
        void* p = nullptr;
        if (int rc = b200ps_ipc_open((const char*)handles + (size_t)r * B200PS_IPC_HANDLE_BYTES, &p)) return rc;
        t->base[r] = (char*)p;
    }
    t->imported = true;
    build_pointers(t);
    return B200PS_OK;
}

extern "C" int b200ps_table_sync(b200ps_table_t* t, void* stream) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    B200PS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return alive(t);
}

extern "C" int b200ps_table_init_uniform(b200ps_table_t* t, uint64_t seed, void* stream) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    for (int s : t->local_shards) {
        const uint64_t b = t->L.shard_range[(size_t)s].first, e = t->L.shard_range[(size_t)s].second;
        if (int rc = b200ps_fill_uniform(t->shard_ptrs[(size_t)s], (e - b) * (uint64_t)t->L.dim, seed, 0, 0, b * (uint64_t)t->L.dim, stream)) return rc;
    }
    return B200PS_OK;
}

extern "C" int b200ps_table_worker_buffers(b200ps_table_t* t, int worker, int64_t** d_ids, float** d_grad_rows, float** d_pulled) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    B200PS_REQUIRE(worker >= 0 && worker < t->L.W && t->L.worker_rank(worker) == t->rank, B200PS_ERR_RANGE,
                   "compute task %d is not placed on rank %d", worker, t->rank);
    if (d_ids) *d_ids = reinterpret_cast<int64_t*>(t->arena + t->L.ids_off[(size_t)worker]);
    if (d_grad_rows) *d_grad_rows = reinterpret_cast<float*>(t->arena + t->L.grad_off[(size_t)worker]);
    if (d_pulled) *d_pulled = reinterpret_cast<float*>(t->arena + t->L.out_off[(size_t)worker]);
    return B200PS_OK;
}

extern "C" int b200ps_table_shard(b200ps_table_t* t, int shard, float** d_rows, float** d_accum, uint64_t* row_begin, uint64_t* row_end) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    B200PS_REQUIRE(shard >= 0 && shard < t->L.S && t->L.shard_rank(shard) == t->rank, B200PS_ERR_RANGE, "shard %d is not placed on rank %d",
                   shard, t->rank);
    if (d_rows) *d_rows = t->shard_ptrs[(size_t)shard];
    if (d_accum) *d_accum = t->accum_ptrs[(size_t)shard];
    if (row_begin) *row_begin = t->L.shard_range[(size_t)shard].first;
    if (row_end) *row_end = t->L.shard_range[(size_t)shard].second;
    return B200PS_OK;
}

extern "C" int b200ps_sparse_pull(b200ps_table_t* t, void* stream) {
    if (int rc = alive(t)) return rc;
    return enqueue_pull(t, (cudaStream_t)stream, true);
}

extern "C" int b200ps_sparse_pull_worker(b200ps_table_t* t, int worker, float* d_out, void* stream) {
    if (int rc = alive(t)) return rc;
    int64_t* ids = nullptr;
    if (int rc = b200ps_table_worker_buffers(t, worker, &ids, nullptr, nullptr)) return rc;
    return b200ps_sparse_gather(d_out, t->shard_cptrs.data(), &t->geom, ids, t->L.B, stream);
}

extern "C" int b200ps_sparse_push(b200ps_table_t* t, void* stream) {
    if (int rc = alive(t)) return rc;
    return enqueue_push(t, (cudaStream_t)stream);
}

// pull of all tasks + push, as one unit (enqueue_step)
extern "C" int b200ps_sparse_step(b200ps_table_t* t, void* stream) {
    if (int rc = alive(t)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (t->o.use_graph && !t->profile) {
        if (!t->graph) {
            cudaGraph_t g = nullptr;
            B200PS_CUDA(cudaStreamBeginCapture(t->s_cap, cudaStreamCaptureModeThreadLocal));
            const int rc = enqueue_step(t, t->s_cap);
            const cudaError_t e = cudaStreamEndCapture(t->s_cap, &g);
            if (rc) {
                if (g) cudaGraphDestroy(g);
                return rc;
            }
            B200PS_CUDA(e);
            size_t n = 0;
            cudaGraphGetNodes(g, nullptr, &n);
            std::vector<cudaGraphNode_t> nodes(n);
            if (n) cudaGraphGetNodes(g, nodes.data(), &n);
            t->graph_kernels = 0;
            for (cudaGraphNode_t nd : nodes) {
                cudaGraphNodeType ty;
                if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel) t->graph_kernels += 1;
            }
            const cudaError_t e2 = cudaGraphInstantiate(&t->graph, g, 0);
            cudaGraphDestroy(g);
            B200PS_CUDA(e2);
        }
        B200PS_CUDA(cudaGraphLaunch(t->graph, st));
        g_launches.fetch_add(t->graph_kernels, std::memory_order_relaxed);
        return B200PS_OK;
    }
    return enqueue_step(t, st);
}

extern "C" int b200ps_table_set_graph(b200ps_table_t* t, int on) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    t->o.use_graph = on != 0;
    return B200PS_OK;
}

extern "C" int b200ps_table_set_profile(b200ps_table_t* t, int on) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    t->profile = on != 0;
    return B200PS_OK;
}

extern "C" int b200ps_table_phase_ms(b200ps_table_t* t, float* serve_ms, float* sort_ms, float* apply_ms) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    B200PS_CUDA(cudaDeviceSynchronize());
    float* out[3] = {serve_ms, sort_ms, apply_ms};
    for (int k = 0; k < 3; ++k) {
        double sum = 0;
        for (const auto& s : t->spans[k]) {
            float ms = 0;
            B200PS_CUDA(cudaEventElapsedTime(&ms, s.a, s.b));
            sum += ms;
        }
        if (out[k]) *out[k] = t->spans[k].empty() ? 0.f : (float)(sum / (double)t->spans[k].size());
        t->spans[k].clear();
    }
    for (cudaEvent_t e : t->event_pool) cudaEventDestroy(e);
    t->event_pool.clear();
    return B200PS_OK;
}

extern "C" int b200ps_table_plan_read(b200ps_table_t* t, uint64_t* n_kept, uint64_t* n_unique, uint64_t* n_bad, int64_t* h_sorted_ids,
                                      uint32_t* h_perm, int32_t* h_seg_id, int64_t* h_unique_ids, uint32_t* h_seg_offset, void* stream) {
    B200PS_REQUIRE(t, B200PS_ERR_INVALID, "null table");
    return b200ps_sparse_plan_read(t->workspace, (uint64_t)t->L.W * t->L.B, t->L.rows, n_kept, n_unique, n_bad, h_sorted_ids, h_perm,
                                   h_seg_id, h_unique_ids, h_seg_offset, stream);
}

extern "C" int b200ps_table_save(b200ps_table_t* t, const char* directory, void* stream) {
    B200PS_REQUIRE(directory, B200PS_ERR_INVALID, "null directory");
    if (int rc = b200ps_table_sync(t, stream)) return rc;
    if (mkdir(directory, 0777) != 0 && errno != EEXIST) return fail(B200PS_ERR_INVALID, "cannot create %s: %s", directory, strerror(errno));
    const std::string path = tbl_path(t, directory), tmp = path + ".tmp" + std::to_string((long)getpid());
    FILE* f = fopen(tmp.c_str(), "wb");
    B200PS_REQUIRE(f, B200PS_ERR_INVALID, "cannot write %s: %s", tmp.c_str(), strerror(errno));
    TblHeader h{};
    memcpy(h.magic, "B200PST2", 8);
    h.rows = t->L.rows;
    h.dim = t->L.dim;
    h.S = t->L.S;
    h.optimizer = t->o.optimizer;
    h.rank = t->rank;
    h.world = t->world;
    h.n_shards = (int32_t)t->local_shards.size();
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    std::vector<float> buf;
    for (int s : t->local_shards) {
        const size_t n = (size_t)(t->L.shard_range[(size_t)s].second - t->L.shard_range[(size_t)s].first) * (size_t)t->L.dim;
        const int32_t id = s;
        ok = ok && fwrite(&id, sizeof id, 1, f) == 1;
        buf.resize(n);
        const float* src[2] = {t->shard_ptrs[(size_t)s], t->L.adagrad ? t->accum_ptrs[(size_t)s] : nullptr};
        for (int k = 0; ok && k < 2; ++k) {
            if (!src[k] || !n) continue;
            ok = cudaMemcpy(buf.data(), src[k], n * 4, cudaMemcpyDeviceToHost) == cudaSuccess && fwrite(buf.data(), 4, n, f) == n;
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

extern "C" int b200ps_table_load(b200ps_table_t* t, const char* directory, void* stream) {
    B200PS_REQUIRE(directory, B200PS_ERR_INVALID, "null directory");
    if (int rc = b200ps_table_sync(t, stream)) return rc;
    const std::string path = tbl_path(t, directory);
    FILE* f = fopen(path.c_str(), "rb");
    B200PS_REQUIRE(f, B200PS_ERR_INVALID, "cannot read %s: %s", path.c_str(), strerror(errno));
    std::unique_ptr<FILE, int (*)(FILE*)> guard(f, fclose);
    TblHeader h{};
    B200PS_REQUIRE(fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, "B200PST2", 8) == 0, B200PS_ERR_INVALID,
                   "%s is not a b200ps table checkpoint (format B200PST2)", path.c_str());
    B200PS_REQUIRE(h.rows == t->L.rows && h.dim == t->L.dim && h.S == t->L.S && h.optimizer == t->o.optimizer && h.world == t->world &&
                       h.rank == t->rank && h.n_shards == (int32_t)t->local_shards.size(),
                   B200PS_ERR_INVALID, "%s was written for a different table geometry (%llu x %d, %d shards, optimizer %d, rank %d of %d)",
                   path.c_str(), (unsigned long long)h.rows, h.dim, h.S, h.optimizer, h.rank, h.world);
    std::vector<float> buf;
    for (int k = 0; k < h.n_shards; ++k) {
        int32_t s = -1;
        B200PS_REQUIRE(fread(&s, sizeof s, 1, f) == 1 && s >= 0 && s < t->L.S && t->L.shard_rank(s) == t->rank, B200PS_ERR_INVALID,
                       "%s is truncated or names a foreign shard", path.c_str());
        const size_t n = (size_t)(t->L.shard_range[(size_t)s].second - t->L.shard_range[(size_t)s].first) * (size_t)t->L.dim;
        buf.resize(n);
        float* dst[2] = {t->shard_ptrs[(size_t)s], t->L.adagrad ? t->accum_ptrs[(size_t)s] : nullptr};
        for (int a = 0; a < 2; ++a) {
            if (!dst[a] || !n) continue;
            B200PS_REQUIRE(fread(buf.data(), 4, n, f) == n, B200PS_ERR_INVALID, "%s is truncated", path.c_str());
            B200PS_CUDA(cudaMemcpy(dst[a], buf.data(), n * 4, cudaMemcpyHostToDevice));
        }
    }
    return B200PS_OK;
}
