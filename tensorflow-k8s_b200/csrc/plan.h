// plan.h -- the per-rank launch plan of the dense PS step (pure host arithmetic; csrc/plan.cpp).
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

#include "b200ps.h"

struct b200ps_plan {
    struct Launch {
        int shard;
        uint64_t begin, end;  // element range of the slice in the flat buffer
        // buffers the kernel folds, in fold order: (rank, arena offset in elements).  FLAT: one per
        // worker, worker order.  TWO_LEVEL: one per GPU that hosts workers (its partial sum), GPU order.
        std::vector<std::pair<int, uint64_t>> src;
        std::vector<int> mirrors;  // ranks whose replica receives the updated slice (all but this one)
    };
    int rank = 0, world = 1, S = 0, W = 0;
    int aggregate = B200PS_AGG_FLAT;
    uint64_t P = 0;  // padded parameter count
    std::vector<std::pair<uint64_t, uint64_t>> slices;  // every shard's range
    std::vector<int> local_workers, local_shards;
    std::vector<Launch> launches;  // one per local shard

    bool two_level() const { return aggregate != B200PS_AGG_FLAT && world > 1; }
    int first_grad_slot() const { return two_level() ? 2 : 1; }  // slot 0 = replica, slot 1 = partial sum (two-level)
    int workers_on(int r) const { return r < W ? (W - r + world - 1) / world : 0; }
    uint64_t arena_elems(int r) const { return P * (uint64_t)(first_grad_slot() + workers_on(r)); }
    std::pair<int, uint64_t> grad_offset(int w) const {
        return {w % world, P * (uint64_t)(first_grad_slot() + w / world)};
    }
};
