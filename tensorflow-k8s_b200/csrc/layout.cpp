// layout.cpp -- PS variable sharding (include/b200ps.h "Layout group").
//
// Reference: none.  The reference only says PS tasks serve parameters (k8s-operator.md:6); how a
// variable set is spread over them is TensorFlow's business inside the user container.  Here the
// variables are concatenated in declaration order into one flat fp32 buffer (each padded to
// `align` elements so every variable and every slice starts on a 4 KB boundary), and the buffer
// is range-partitioned: slice s = [s*c, min((s+1)*c, P_pad)), c = ceil(P_pad / S) rounded up to
// `align`.  Embedding rows use TensorFlow's "div" partitioning.
#include <memory>
#include <vector>

#include "common.h"

struct b200ps_shardmap {
    std::vector<uint64_t> var_offset;
    std::vector<uint64_t> begin, end;
    uint64_t total = 0;
};

static uint64_t round_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

extern "C" int b200ps_shard_map_build(const uint64_t* var_numel, int nvars, int nshards, uint32_t align,
                                      b200ps_shardmap_t** out) {
    using namespace b200ps;
    B200PS_REQUIRE(out && (var_numel || nvars == 0), B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(nvars >= 0 && nshards >= 1 && align >= 1, B200PS_ERR_INVALID,
                   "shard map needs nvars >= 0, nshards >= 1, align >= 1");
    auto m = std::make_unique<b200ps_shardmap>();
    uint64_t off = 0;
    for (int v = 0; v < nvars; ++v) {
        m->var_offset.push_back(off);
        B200PS_REQUIRE(var_numel[v] <= (UINT64_MAX >> 4) - off, B200PS_ERR_RANGE, "variable %d overflows the layout", v);
        off += round_up(var_numel[v], align);
    }
    m->total = off;
    const uint64_t c = round_up((off + (uint64_t)nshards - 1) / (uint64_t)nshards, align);
    for (int s = 0; s < nshards; ++s) {
        uint64_t b = (uint64_t)s * c, e = b + c;
        m->begin.push_back(b < off ? b : off);
        m->end.push_back(e < off ? e : off);
    }
    *out = m.release();
    return B200PS_OK;
}

extern "C" void b200ps_shard_map_destroy(b200ps_shardmap_t* m) { delete m; }
extern "C" uint64_t b200ps_shard_map_total(const b200ps_shardmap_t* m) { return m ? m->total : 0; }
extern "C" int b200ps_shard_map_nshards(const b200ps_shardmap_t* m) { return m ? (int)m->begin.size() : B200PS_ERR_INVALID; }
extern "C" int b200ps_shard_map_nvars(const b200ps_shardmap_t* m) { return m ? (int)m->var_offset.size() : B200PS_ERR_INVALID; }

extern "C" int b200ps_shard_map_slice(const b200ps_shardmap_t* m, int shard, uint64_t* begin, uint64_t* end) {
    using namespace b200ps;
    B200PS_REQUIRE(m && begin && end, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(shard >= 0 && shard < (int)m->begin.size(), B200PS_ERR_RANGE, "no shard %d", shard);
    *begin = m->begin[(size_t)shard];
    *end = m->end[(size_t)shard];
    return B200PS_OK;
}

extern "C" int b200ps_shard_map_var_offset(const b200ps_shardmap_t* m, int var, uint64_t* offset) {
    using namespace b200ps;
    B200PS_REQUIRE(m && offset, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(var >= 0 && var < (int)m->var_offset.size(), B200PS_ERR_RANGE, "no variable %d", var);
    *offset = m->var_offset[(size_t)var];
    return B200PS_OK;
}

extern "C" int b200ps_row_range(uint64_t rows, int nshards, int shard, uint64_t* begin, uint64_t* end) {
    using namespace b200ps;
    B200PS_REQUIRE(begin && end, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(nshards >= 1 && shard >= 0 && shard < nshards, B200PS_ERR_RANGE, "no shard %d of %d", shard, nshards);
    const uint64_t q = rows / (uint64_t)nshards, r = rows % (uint64_t)nshards, s = (uint64_t)shard;
    *begin = s * q + (s < r ? s : r);  // the first r shards hold one row more
    *end = *begin + q + (s < r ? 1 : 0);
    return B200PS_OK;
}

extern "C" int b200ps_row_owner(uint64_t rows, int nshards, uint64_t id) {
    using namespace b200ps;
    B200PS_REQUIRE(nshards >= 1, B200PS_ERR_INVALID, "nshards must be >= 1");
    B200PS_REQUIRE(id < rows, B200PS_ERR_RANGE, "row id %llu is outside the table of %llu rows", (unsigned long long)id,
                   (unsigned long long)rows);
    const uint64_t q = rows / (uint64_t)nshards, r = rows % (uint64_t)nshards, t = r * (q + 1);
    if (id < t) return (int)(id / (q + 1));
    return (int)(r + (id - t) / q);
}
