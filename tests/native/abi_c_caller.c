/* A plain C99 caller of libb200ps.so -- what a cgo binding would do: TFJob text -> cluster ->
 * shard map -> context -> two Adam steps with synthetic gradients -> dump.  No Python, no torch:
 * only the C ABI of include/b200ps.h and the CUDA runtime for nothing (the library owns all device
 * memory).  Writes the dumped parameters / m / v to a file the test compares with the oracle.
 *
 *   abi_c_caller <out.bin> [graph]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200ps.h"

#define CHECK(call)                                                                   \
    do {                                                                              \
        int rc_ = (call);                                                             \
        if (rc_ < 0) {                                                                \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc_, b200ps_last_error());       \
            return 1;                                                                 \
        }                                                                             \
    } while (0)

static const char* TFJOB =
    "{\"apiVersion\":\"tensorflow.org/v1alpha1\",\"kind\":\"TFJob\",\"metadata\":{\"name\":\"c-caller\"},"
    "\"spec\":{\"replicaSpecs\":[{\"replicas\":2,\"tfReplicaType\":\"PS\"},{\"replicas\":3,\"tfReplicaType\":\"WORKER\"}]}}";

int main(int argc, char** argv) {
    const uint64_t vars[3] = {70000, 300, 33000};
    b200ps_cluster_t* cluster = NULL;
    b200ps_shardmap_t* map = NULL;
    b200ps_ctx_t* ctx = NULL;
    b200ps_dense_options_t opt;
    char tf_config[2048];
    uint64_t P, t;
    int W, w, step;
    float *p, *m, *v;
    FILE* f;

    if (argc < 2) return 2;
    CHECK(b200ps_cluster_from_tfjob(TFJOB, &cluster));
    CHECK(b200ps_tf_config_for(cluster, "worker", 1, tf_config, sizeof tf_config));
    printf("TF_CONFIG %s\n", tf_config);
    W = b200ps_cluster_num_workers(cluster);
    CHECK(b200ps_shard_map_build(vars, 3, b200ps_cluster_num_shards(cluster), B200PS_ALIGN_ELEMS, &map));
    CHECK(b200ps_dense_options_default(&opt));
    opt.optimizer = B200PS_OPT_ADAM;
    opt.lr = 1e-3;
    opt.use_graph = argc > 2 && strcmp(argv[2], "graph") == 0;
    CHECK(b200ps_ctx_create(cluster, map, 0, 1, &opt, &ctx));
    P = b200ps_plan_params_padded(b200ps_ctx_plan(ctx));
    CHECK(b200ps_dense_init_uniform(ctx, 1, NULL));
    for (step = 1; step <= 2; ++step) {
        for (w = 0; w < W; ++w) { /* the compute tasks "compute": synthetic gradients straight into the arena */
            float* g = NULL;
            CHECK(b200ps_dense_grad_buffer(ctx, w, &g));
            CHECK(b200ps_fill_uniform(g, P, 2, (uint64_t)step, (uint64_t)w, 0, NULL));
        }
        CHECK(b200ps_dense_step(ctx, NULL));
    }
    t = b200ps_ctx_step_count(ctx);
    p = (float*)malloc(P * 4);
    m = (float*)calloc(P, 4);
    v = (float*)calloc(P, 4);
    CHECK(b200ps_dense_dump(ctx, p, m, v, NULL));
    f = fopen(argv[1], "wb");
    if (!f) return 3;
    fwrite(&P, 8, 1, f);
    fwrite(&t, 8, 1, f);
    fwrite(p, 4, P, f);
    fwrite(m, 4, P, f);
    fwrite(v, 4, P, f);
    fclose(f);
    printf("P_pad %llu steps %llu launches %llu\n", (unsigned long long)P, (unsigned long long)t,
           (unsigned long long)b200ps_launch_count());
    CHECK(b200ps_ctx_destroy(ctx));
    b200ps_shard_map_destroy(map);
    b200ps_cluster_destroy(cluster);
    free(p);
    free(m);
    free(v);
    return 0;
}
