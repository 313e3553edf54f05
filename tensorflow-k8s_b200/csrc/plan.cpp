// plan.cpp -- who folds what from where: the launch plan of one rank (include/b200ps.h "Context
// group", b200ps_plan_*).  Pure host arithmetic, so world_size > 1 can be tested on a CPU.
//
// Reference: none -- in a TFJob the placement of variables on PS tasks and the gradient routes
// between WORKER and PS tasks (k8s-operator.md:6) are TensorFlow's device placement and gRPC
// rendezvous.  Here: shard s -> GPU s mod G, compute task w -> GPU w mod G (b200ps_placement_gpu);
// a rank's arena is [replica | (two-level: partial sum) | one gradient buffer per local worker];
// the kernel of a local shard folds, in fixed order, either every worker's gradient slice (FLAT)
// or one partial sum per GPU that hosts workers (TWO_LEVEL / MULTIMEM), and mirrors the updated
// slice into every other rank's replica.
#include "plan.h"

#include <memory>

#include "common.h"

extern "C" int b200ps_plan_build(const b200ps_shardmap_t* map, int num_workers, int rank, int world, int aggregate,
                                 b200ps_plan_t** out) {
    using namespace b200ps;
    B200PS_REQUIRE(map && out, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(aggregate == B200PS_AGG_FLAT || aggregate == B200PS_AGG_TWO_LEVEL || aggregate == B200PS_AGG_MULTIMEM,
                   B200PS_ERR_INVALID, "aggregate must be B200PS_AGG_FLAT, _TWO_LEVEL or _MULTIMEM");
    B200PS_REQUIRE(world >= 1 && rank >= 0 && rank < world, B200PS_ERR_RANGE, "rank %d is outside 0..%d", rank, world - 1);
    B200PS_REQUIRE(world - 1 <= B200PS_MAX_MIRRORS, B200PS_ERR_RANGE,
                   "one launch writes at most %d peer replicas (world <= %d)", B200PS_MAX_MIRRORS, B200PS_MAX_MIRRORS + 1);
    B200PS_REQUIRE(num_workers >= 1, B200PS_ERR_RANGE, "a job needs at least one compute task");
    auto p = std::make_unique<b200ps_plan>();
    p->rank = rank;
    p->world = world;
    p->S = b200ps_shard_map_nshards(map);
    p->W = num_workers;
    p->aggregate = aggregate;
    p->P = b200ps_shard_map_total(map);
    for (int s = 0; s < p->S; ++s) {
        uint64_t b = 0, e = 0;
        if (int rc = b200ps_shard_map_slice(map, s, &b, &e)) return rc;
        p->slices.emplace_back(b, e);
    }
    for (int w = 0; w < num_workers; ++w)
        if (b200ps_placement_gpu(w, world) == rank) p->local_workers.push_back(w);
    for (int s = 0; s < p->S; ++s)
        if (b200ps_placement_gpu(s, world) == rank) p->local_shards.push_back(s);
    std::vector<std::pair<int, uint64_t>> src;
    if (p->two_level()) {
        for (int r = 0; r < world; ++r)
            if (p->workers_on(r) > 0) src.emplace_back(r, p->P);  // slot 1 of rank r's arena: its partial sum
    } else {
        for (int w = 0; w < num_workers; ++w) src.push_back(p->grad_offset(w));
    }
    B200PS_REQUIRE((int)src.size() <= B200PS_MAX_WORKERS, B200PS_ERR_RANGE, "one launch folds at most %d buffers (got %zu)",
                   B200PS_MAX_WORKERS, src.size());
    for (int s : p->local_shards) {
        b200ps_plan::Launch L;
        L.shard = s;
        L.begin = p->slices[(size_t)s].first;
        L.end = p->slices[(size_t)s].second;
        L.src = src;
        for (int r = 0; r < world; ++r)
            if (r != rank) L.mirrors.push_back(r);
        p->launches.push_back(std::move(L));
    }
    *out = p.release();
    return B200PS_OK;
}

extern "C" void b200ps_plan_destroy(b200ps_plan_t* p) { delete p; }
extern "C" uint64_t b200ps_plan_params_padded(const b200ps_plan_t* p) { return p ? p->P : 0; }
extern "C" int b200ps_plan_two_level(const b200ps_plan_t* p) { return p ? (p->two_level() ? 1 : 0) : B200PS_ERR_INVALID; }
extern "C" int b200ps_plan_first_grad_slot(const b200ps_plan_t* p) { return p ? p->first_grad_slot() : B200PS_ERR_INVALID; }
extern "C" int b200ps_plan_num_local_workers(const b200ps_plan_t* p) { return p ? (int)p->local_workers.size() : B200PS_ERR_INVALID; }
extern "C" int b200ps_plan_num_launches(const b200ps_plan_t* p) { return p ? (int)p->launches.size() : B200PS_ERR_INVALID; }

extern "C" uint64_t b200ps_plan_arena_elems(const b200ps_plan_t* p, int of_rank) {
    return (p && of_rank >= 0 && of_rank < p->world) ? p->arena_elems(of_rank) : 0;
}

extern "C" int b200ps_plan_local_worker(const b200ps_plan_t* p, int i) {
    using namespace b200ps;
    B200PS_REQUIRE(p, B200PS_ERR_INVALID, "null plan");
    B200PS_REQUIRE(i >= 0 && i < (int)p->local_workers.size(), B200PS_ERR_RANGE, "no local worker %d", i);
    return p->local_workers[(size_t)i];
}

extern "C" int b200ps_plan_grad_offset(const b200ps_plan_t* p, int worker, int* rank, uint64_t* offset) {
    using namespace b200ps;
    B200PS_REQUIRE(p && rank && offset, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(worker >= 0 && worker < p->W, B200PS_ERR_RANGE, "no compute task %d", worker);
    const auto g = p->grad_offset(worker);
    *rank = g.first;
    *offset = g.second;
    return B200PS_OK;
}

extern "C" int b200ps_plan_launch(const b200ps_plan_t* p, int i, int* shard, uint64_t* begin, uint64_t* end, int* n_src,
                                  int* src_rank, uint64_t* src_offset, int* n_mirrors, int* mirror_rank) {
    using namespace b200ps;
    B200PS_REQUIRE(p, B200PS_ERR_INVALID, "null plan");
    B200PS_REQUIRE(i >= 0 && i < (int)p->launches.size(), B200PS_ERR_RANGE, "no launch %d", i);
    const auto& L = p->launches[(size_t)i];
    if (shard) *shard = L.shard;
    if (begin) *begin = L.begin;
    if (end) *end = L.end;
    if (n_src) *n_src = (int)L.src.size();
    for (size_t k = 0; k < L.src.size(); ++k) {
        if (src_rank) src_rank[k] = L.src[k].first;
        if (src_offset) src_offset[k] = L.src[k].second;
    }
    if (n_mirrors) *n_mirrors = (int)L.mirrors.size();
    for (size_t k = 0; k < L.mirrors.size(); ++k)
        if (mirror_rank) mirror_rank[k] = L.mirrors[k];
    return B200PS_OK;
}
