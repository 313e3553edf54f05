// Drives the host-side C++ of libb200ps (contract.cpp, layout.cpp, common.cpp) under
// AddressSanitizer + UBSan: well-formed inputs, every malformed JSON shape we could think of, and
// a few thousand random mutations.  Built and run by tests/test_host_sanitizers.py.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "b200ps.h"

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd() {
    rng_state ^= rng_state << 13;
    rng_state ^= rng_state >> 7;
    rng_state ^= rng_state << 17;
    return rng_state;
}

static int try_tfjob(const std::string& s) {
    b200ps_cluster_t* c = nullptr;
    int rc = b200ps_cluster_from_tfjob(s.c_str(), &c);
    if (rc == 0) {
        char buf[1 << 16];
        for (const char* job : {"ps", "worker", "master", "chief", "nope"}) {
            int n = b200ps_cluster_num_tasks(c, job);
            for (int i = -1; i <= n; ++i) {
                b200ps_tf_config_for(c, job, i, buf, sizeof buf);
                b200ps_tf_config_for(c, job, i, buf, 8);  // too small on purpose
                b200ps_cluster_task_address(c, job, i, buf, sizeof buf);
                b200ps_cluster_worker_rank(c, job, i);
            }
        }
        b200ps_cluster_num_shards(c);
        b200ps_cluster_num_workers(c);
        b200ps_cluster_destroy(c);
    } else if (b200ps_last_error()[0] == 0) {
        fprintf(stderr, "failure without a message for: %s\n", s.c_str());
        return 1;
    }
    return 0;
}

static int try_tf_config(const std::string& s) {
    b200ps_cluster_t* c = nullptr;
    char type[8];
    int idx = -1;
    int rc = b200ps_cluster_from_tf_config(s.c_str(), &c, type, sizeof type, &idx);
    if (rc == 0) {
        char buf[4096];
        b200ps_tf_config_for(c, type, idx, buf, sizeof buf);
        b200ps_cluster_destroy(c);
    }
    return 0;
}

int main() {
    const std::string good =
        "{\"metadata\":{\"name\":\"mlp\"},\"spec\":{\"runtimeId\":\"ab12\",\"replicaSpecs\":[{\"tfReplicaType\":\"MASTER\"},"
        "{\"replicas\":2,\"tfReplicaType\":\"PS\",\"tfPort\":3333},{\"replicas\":4,\"tfReplicaType\":\"WORKER\"}]}}";
    const std::string good_cfg =
        "{\"cluster\":{\"ps\":[\"a:1\"],\"worker\":[\"b:1\",\"c\\u00e9\\ud83d\\ude00:1\"]},\"task\":{\"type\":\"worker\",\"index\":1}}";
    int bad = 0;
    bad += try_tfjob(good);
    bad += try_tf_config(good_cfg);
    const char* shapes[] = {"", "{", "}", "[", "[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[[", "{\"a\":",
                            "{\"a\":1,}", "{\"spec\":{\"replicaSpecs\":[{\"replicas\":1e99}]}}", "{\"spec\":{\"replicaSpecs\":[{\"replicas\":-1}]}}",
                            "{\"spec\":{\"replicaSpecs\":[{\"tfReplicaType\":7}]}}", "{\"spec\":{\"tfReplicaSpecs\":{\"PS\":3}}}", "\"\\u12\"",
                            "{\"metadata\":{\"name\":\"\\ud800\"},\"spec\":{\"replicaSpecs\":[{}]}}", "nul", "tru", "-", "1e", "{\"a\":\"\\", "{\"\":{}}"};
    for (const char* s : shapes) {
        bad += try_tfjob(s);
        bad += try_tf_config(s);
    }
    for (int it = 0; it < 4000; ++it) {  // random byte mutations, truncations and splices of the good inputs
        std::string s = (it & 1) ? good : good_cfg;
        const int edits = 1 + (int)(rnd() % 4);
        for (int e = 0; e < edits && !s.empty(); ++e) {
            const size_t pos = rnd() % s.size();
            switch (rnd() % 4) {
                case 0: s[pos] = (char)(rnd() % 256); break;
                case 1: s.erase(pos, 1 + rnd() % 5); break;
                case 2: s.insert(pos, 1, "{}[]\",:\\u0"[rnd() % 11]); break;
                default: s.resize(pos); break;
            }
        }
        for (char& ch : s)
            if (ch == 0) ch = ' ';
        bad += (it & 1) ? try_tfjob(s) : try_tf_config(s);
    }
    // layout: sizes at the edges
    for (int it = 0; it < 2000; ++it) {
        const int nv = (int)(rnd() % 6), S = 1 + (int)(rnd() % 9);
        std::vector<uint64_t> v((size_t)nv);
        for (auto& x : v) x = (rnd() % 8 == 0) ? (UINT64_MAX >> (rnd() % 8)) : rnd() % 100000;
        b200ps_shardmap_t* m = nullptr;
        if (b200ps_shard_map_build(v.data(), nv, S, 1u << (rnd() % 12), &m) == 0) {
            uint64_t b, e, prev = 0;
            for (int s = 0; s < S; ++s) {
                if (b200ps_shard_map_slice(m, s, &b, &e) != 0 || b != prev || e < b) { fprintf(stderr, "bad slice\n"); ++bad; }
                prev = e;
            }
            if (prev != b200ps_shard_map_total(m)) { fprintf(stderr, "slices do not cover\n"); ++bad; }
            b200ps_shard_map_slice(m, S, &b, &e);
            b200ps_shard_map_var_offset(m, nv, &b);
            b200ps_shard_map_destroy(m);
        }
        const uint64_t rows = 1 + rnd() % 100000;
        const uint64_t id = rnd() % (rows + 2);
        const int o = b200ps_row_owner(rows, S, id);
        if (id < rows) {
            uint64_t rb, re;
            if (o < 0 || b200ps_row_range(rows, S, o, &rb, &re) != 0 || id < rb || id >= re) { fprintf(stderr, "owner/range disagree\n"); ++bad; }
        }
    }
    printf("host_fuzz: %d problems\n", bad);
    return bad ? 1 : 0;
}
