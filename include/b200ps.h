/*
 * b200ps.h -- C ABI of libb200ps.so: the parameter-server (PS) step of a TFJob on one B200 box.
 *
 * What this boundary replaces.  The reference (yph152/tensorflow-k8s) is the notes of a TFJob
 * operator; it has NO FFI, plugin or kernel interface for the PS data path -- the path lives inside
 * the TensorFlow containers the operator launches.  Its only statement about the path is the role
 * model at k8s-operator.md:6 (PS tasks serve parameters, WORKER tasks compute, together a cluster);
 * the only files that own the operator-facing contract are named, not shown, in images/tf3.PNG
 * (pkg/apis/tensorflow/v1alpha1/{types,defaults}.go, pkg/apis/tensorflow/validation/validation.go,
 * pkg/trainer/replicas.go).  Each group below cites the item it stands in for.  A Go operator
 * would bind these through cgo (INTEGRATION.md shows the stub).
 *
 * Conventions
 *   - plain C99: pointers, sizes, opaque handles; no C++ or torch types.
 *   - every function returns 0 on success or a negative b200ps_status; the message for the last
 *     failure on the calling thread is b200ps_last_error().  No abort(), no exit().
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Kernel
 *     entry points are asynchronous and stream-ordered; device buffers passed in must stay alive
 *     until the stream reaches that point.
 *   - pointers named d_* are device pointers on the current device (or peer-mapped pointers of
 *     another GPU on the same NVSwitch domain); h_* are host pointers.
 *   - there is no CPU fallback: without a CUDA device the kernel entry points fail.
 */
#ifndef B200PS_H
#define B200PS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PS_MAX_WORKERS 64 /* gradient buffers one launch can fold */
#define B200PS_MAX_MIRRORS 8  /* extra (peer) copies of the updated slice one launch can write */
#define B200PS_ALIGN_ELEMS 1024u /* default shard-map alignment, in fp32 elements */

typedef enum {
    B200PS_OK = 0,
    B200PS_ERR_INVALID = -1, /* bad argument */
    B200PS_ERR_PARSE = -2,   /* malformed TFJob / TF_CONFIG JSON */
    B200PS_ERR_VALIDATION = -3, /* TFJob spec rejected by validation */
    B200PS_ERR_CUDA = -4,    /* CUDA runtime error (message holds cudaGetErrorString) */
    B200PS_ERR_RANGE = -5,   /* index / size out of range, buffer too small */
    B200PS_ERR_STATE = -6    /* call order violated */
} b200ps_status;

const char* b200ps_last_error(void);
int b200ps_version(void);

/* ------------------------------------------------------------------------------------------------
 * Contract group: TFJob replica specs in, TF_CONFIG per task out.
 * Stands in for pkg/apis/tensorflow/v1alpha1/{types,defaults}.go + validation/validation.go
 * (spec, defaults replicas=1 / tfPort=2222, validation) and pkg/trainer/replicas.go (TF_CONFIG
 * generation) -- file names in images/tf3.PNG; roles per k8s-operator.md:6.
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200ps_cluster b200ps_cluster_t;

/* tfjob_json: a TFJob object as JSON (the launcher converts YAML first).  Accepts the v1alpha1
 * list form spec.replicaSpecs[{replicas,tfReplicaType,tfPort}] (types MASTER|PS|WORKER) and the
 * later map form spec.tfReplicaSpecs{Chief|Master|PS|Worker:{replicas,...}}. */
int b200ps_cluster_from_tfjob(const char* tfjob_json, b200ps_cluster_t** out);
/* Parses a TF_CONFIG value; task_type receives "ps" | "worker" | "master" | "chief". */
int b200ps_cluster_from_tf_config(const char* tf_config_json, b200ps_cluster_t** out,
                                  char* task_type, size_t task_type_cap, int* task_index);
void b200ps_cluster_destroy(b200ps_cluster_t* c);
/* number of tasks of a job ("ps", "worker", "master", "chief"); 0 if the job is absent */
int b200ps_cluster_num_tasks(const b200ps_cluster_t* c, const char* job);
/* writes "host:port" of one task; returns its length or a negative status */
int b200ps_cluster_task_address(const b200ps_cluster_t* c, const char* job, int index, char* buf,
                                size_t cap);
/* Exact TF_CONFIG string for one task:
 * {"cluster":{"master":[..],"ps":[..],"worker":[..]},"task":{"type":"worker","index":0},"environment":"cloud"}
 * (cluster keys in sorted order, no whitespace).  Returns the length or a negative status. */
int b200ps_tf_config_for(const b200ps_cluster_t* c, const char* job, int index, char* buf,
                         size_t cap);
/* Data-path view of the cluster: S = #ps tasks; W = #compute tasks (master/chief first, then
 * workers -- the master/chief trains too). */
int b200ps_cluster_num_shards(const b200ps_cluster_t* c);
int b200ps_cluster_num_workers(const b200ps_cluster_t* c);
/* compute-task rank of (job, index) in [0, W), or negative if it is not a compute task */
int b200ps_cluster_worker_rank(const b200ps_cluster_t* c, const char* job, int index);
/* single-box placement: ps shard s -> GPU s mod G, compute task w -> GPU w mod G */
int b200ps_placement_gpu(int task_rank, int ngpus);

/* ------------------------------------------------------------------------------------------------
 * Layout group: PS variable sharding (range partitioning of the flat parameter buffer; TF "div"
 * partitioning of embedding rows).  Reference: none -- TensorFlow's partitioners inside the user
 * container; roles per k8s-operator.md:6.  SURVEY.md 8(a) row a4.
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200ps_shardmap b200ps_shardmap_t;

int b200ps_shard_map_build(const uint64_t* var_numel, int nvars, int nshards, uint32_t align,
                           b200ps_shardmap_t** out);
void b200ps_shard_map_destroy(b200ps_shardmap_t* m);
uint64_t b200ps_shard_map_total(const b200ps_shardmap_t* m); /* P_pad */
int b200ps_shard_map_nshards(const b200ps_shardmap_t* m);
int b200ps_shard_map_nvars(const b200ps_shardmap_t* m);
int b200ps_shard_map_slice(const b200ps_shardmap_t* m, int shard, uint64_t* begin, uint64_t* end);
int b200ps_shard_map_var_offset(const b200ps_shardmap_t* m, int var, uint64_t* offset);
int b200ps_row_range(uint64_t rows, int nshards, int shard, uint64_t* begin, uint64_t* end);
int b200ps_row_owner(uint64_t rows, int nshards, uint64_t id); /* owner shard or negative */

/* ------------------------------------------------------------------------------------------------
 * Kernel group (stateless; device pointers; sm_100a).  Reference: none -- TensorFlow's
 * ApplyGradientDescent / ApplyAdam CPU ops on the PS task and its gRPC tensor transport;
 * k8s-operator.md:6.  SURVEY.md 8(a) rows a5..a9.
 *
 * One launch is push + aggregate + apply + pull for one PS slice of n elements:
 *   g[i]  = inv_w * (((g_0[i] + g_1[i]) + g_2[i]) + ...)         fixed buffer order, W buffers
 *   SGD:    p[i] -= lr * g[i]
 *   Adam:   m = b1*m + omb1*g; v = b2*v + omb2*g*g; p -= (lr_t*m) / (sqrt(v) + eps)
 * then the new p[i] is also written to every d_mirrors[k][i] (the workers' copies on other
 * GPUs).  d_grads[w] and d_mirrors[k] may be peer-mapped pointers: the loads are the push over
 * NVLink, the stores the pull.  All arithmetic is fp32 with one IEEE rounding per operation (no
 * FMA contraction), bit-identical to oracle/ps_oracle.c.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    float lr_t; /* lr * sqrt(1 - beta2^t) / (1 - beta1^t) */
    float beta1, beta2;
    float omb1, omb2; /* (float)(1.0 - beta) */
    float eps;
} b200ps_adam_t;

/* bias-corrected scalars for step t >= 1, evaluated in double and rounded once */
int b200ps_adam_scalars(double lr, double beta1, double beta2, double eps, uint64_t t,
                        b200ps_adam_t* out);

/* inv_w is the fp32 constant the folded sum is multiplied by: 1/W for the plain mean over W
 * workers; 1/(total workers) when the W buffers are themselves per-GPU partial sums (below). */
int b200ps_dense_apply_sgd(float* d_params, const float* const* d_grads, int W, uint64_t n,
                           float inv_w, float lr, float* const* d_mirrors, int n_mirrors,
                           void* stream);
int b200ps_dense_apply_adam(float* d_params, float* d_m, float* d_v, const float* const* d_grads,
                            int W, uint64_t n, float inv_w, const b200ps_adam_t* opt,
                            float* const* d_mirrors, int n_mirrors, void* stream);
/* TensorFlow's ApplyMomentum (no Nesterov): accum = momentum*accum + g; p -= lr*accum.
 * (more optimizers.) */
int b200ps_dense_apply_momentum(float* d_params, float* d_accum, const float* const* d_grads, int W,
                                uint64_t n, float inv_w, float lr, float momentum,
                                float* const* d_mirrors, int n_mirrors, void* stream);
/* TensorFlow's ApplyFtrl with lr_power = -0.5 (the "wide" part of Wide&Deep):
 *   na = accum + g*g;  sigma = (sqrt(na) - sqrt(accum)) / lr;  linear += g - sigma*p;
 *   p = |linear| > l1 ? (sign(linear)*l1 - linear) / (sqrt(na)/lr + 2*l2) : 0;  accum = na
 * accum starts at 0.1 (TF's initial_accumulator_value), linear at 0. */
int b200ps_dense_apply_ftrl(float* d_params, float* d_accum, float* d_linear, const float* const* d_grads,
                            int W, uint64_t n, float inv_w, float lr, float l1, float l2,
                            float* const* d_mirrors, int n_mirrors, void* stream);
/* d_out[i] = ((b_0[i] + b_1[i]) + b_2[i]) + ...  -- the local half of the two-level aggregation:
 * a GPU that hosts several workers folds their gradients once (in worker order), and only that
 * partial sum crosses NVLink; the owning shard then folds the G partials in GPU order. */
int b200ps_dense_fold(float* d_out, const float* const* d_bufs, int W, uint64_t n, void* stream);
/* synthetic inputs: d_dst[i] = uniform[-1,1) from the integer hash of (seed, step, worker,
 * first+i); identical to ps_fill_uniform of the oracle */
int b200ps_fill_uniform(float* d_dst, uint64_t n, uint64_t seed, uint64_t step, uint64_t worker,
                        uint64_t first, void* stream);
/* launch geometry of the dense kernels: threads per CTA (128..1024), CTAs per SM, float4 per
 * thread per trip (1, 2 or 4).  0 keeps the current value.  Process-wide; for tuning sweeps. */
int b200ps_set_dense_tuning(int threads, int ctas_per_sm, int unroll);
/* number of kernel launches this library has issued in the calling process */
uint64_t b200ps_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Worker group (SURVEY.md 8(f) rank 1 -- the step on the other side of push / pull).  Reference:
 * none -- the user's TensorFlow program in the WORKER containers (k8s-operator.md:6, "WORKER
 * computes").  A real worker for BASELINE configs[0]: the mini-batch gradient of a linear
 * regression on synthetic data generated in the kernels (x_b[i] = u(seed_x, step, worker*B + b, i),
 * y_b = x_b . w_true, w_true[i] = u(seed_w, 0, 0, i)); fixed-order double accumulation, one rounding
 * to float, bit-identical to ps_linreg_grad of the oracle.  d_grad is typically a gradient buffer
 * of the PS arena (zero-copy push), d_params its replica (zero-copy pull).
 * ---------------------------------------------------------------------------------------------- */
size_t b200ps_linreg_scratch_bytes(uint64_t n, int batch);
int b200ps_linreg_grad(float* d_grad, const float* d_params, uint64_t n, int batch, uint64_t seed_x,
                       uint64_t seed_w, uint64_t step, uint64_t worker, void* d_scratch, size_t scratch_bytes,
                       void* stream);

/* ------------------------------------------------------------------------------------------------
 * MLP worker group: the compute tasks of BASELINE configs[1] on the GPU.
 * Reference: none -- the user's TensorFlow program in the WORKER containers (k8s-operator.md:6).
 * Forward and backward of a dense MLP (ReLU between layers, linear output, loss 1/(2B) sum ||y - t||^2)
 * for ALL compute tasks placed on this GPU; each task's fp32 gradient is written straight into its
 * gradient buffer of the PS arena (zero-copy push), the fp32 parameters are read from the replica
 * (zero-copy pull).  Mixed precision: bf16 operands, fp32 accumulation on the tcgen05 tensor cores
 * (csrc/gemm_sm100.cuh: TMA loads, TMEM accumulators), fp32 gradients.  The variables are laid out
 * as synth.mlp_c1_variables does: W_0 [d_0][d_1], b_0 [d_1], W_1, b_1, ...
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200ps_mlp b200ps_mlp_t;
/* dims: d_0 .. d_L (multiples of 128); var_offsets: 2 L offsets (W_l, b_l) into the flat buffers, as
 * b200ps_shard_map_var_offset gives them; batch_per_task a multiple of 64; 1..8 local tasks */
int b200ps_mlp_create(const int* dims, int n_dims, int batch_per_task, int local_tasks, const uint64_t* var_offsets,
                      b200ps_mlp_t** out);
int b200ps_mlp_destroy(b200ps_mlp_t* mlp);
/* One training step of the local tasks: h_x = features bf16 [tasks * batch][d_0] and h_t = targets bf16
 * [tasks * batch][d_L] in pinned host memory (the mini-batches; copied to the device on `stream`),
 * d_params = the replica, d_grads[i] = gradient buffer of the i-th local task; h_loss (may be NULL)
 * receives one loss per task (device -> host on `stream`).  Asynchronous. */
int b200ps_mlp_step(b200ps_mlp_t* mlp, const float* d_params, float* const* d_grads, const void* h_x, const void* h_t,
                    float* h_loss, void* stream);
/* The worker's GEMM on its own, for tests and roofline measurements: C[M][N] (bf16) = A[M][K] * B[N][K]^T,
 * bf16 row-major device buffers, fp32 accumulation; d_ct (may be NULL) receives C^T.  M, N multiples of
 * 128, K a multiple of 64. */
int b200ps_gemm_bf16(const void* d_a, const void* d_b, void* d_c, void* d_ct, int M, int N, int K, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Transport group: buffers other ranks of the box can address.  Reference: none -- TensorFlow's
 * gRPC tensor transport between the PS and WORKER tasks of k8s-operator.md:6.  One process per
 * GPU; each rank exports its gradient / parameter arena, every peer maps it, and the step kernel
 * reads gradients from and writes parameters to the mapped pointers over NVLink (push and pull
 * fused into the apply).  The handle bytes travel over any host channel (the launcher uses
 * torch.distributed).
 * ---------------------------------------------------------------------------------------------- */
#define B200PS_IPC_HANDLE_BYTES 64
int b200ps_dev_alloc(size_t bytes, void** d_ptr); /* cudaMalloc on the current device */
int b200ps_dev_free(void* d_ptr);
int b200ps_ipc_export(const void* d_ptr, void* handle_out /* B200PS_IPC_HANDLE_BYTES */);
int b200ps_ipc_open(const void* handle, void** d_ptr); /* maps a peer's allocation */
int b200ps_ipc_close(void* d_ptr);
/* Stream-ordered barrier across the ranks through peer memory (a few microseconds; the library
 * all-reduce it replaces costs tens).  Every rank owns B200PS_BARRIER_WORDS zero-initialised
 * 32-bit words inside its exported arena; d_flags_by_rank[r] points at rank r's words.  All ranks
 * call it with the same, strictly increasing `epoch` (1, 2, 3, ...).  Work queued on `stream`
 * before the call on every rank is complete -- peer stores included -- before work queued after
 * it on any rank starts.  The wait is bounded (~4 s); a timeout is reported by
 * b200ps_peer_barrier_error (which synchronises the stream). */
#define B200PS_BARRIER_WORDS 16
int b200ps_peer_barrier(uint32_t* const* d_flags_by_rank, int rank, int world, uint32_t epoch, void* stream);
int b200ps_peer_barrier_error(const uint32_t* d_my_flags, void* stream);
/* One-way signals for pipelining a step slice by slice.  signal: after the work queued before it
 * on `stream` has completed, store `epoch` into one flag word (a peer pointer into the consumer's
 * arena).  wait: hold `stream` until d_flags[r] >= epoch for every rank r set in rank_mask (words
 * in local memory; bounded like the barrier, a timeout sets *d_error_word = 1). */
int b200ps_peer_signal(uint32_t* d_flag, uint32_t epoch, void* stream);
int b200ps_peer_wait(const uint32_t* d_flags, uint32_t rank_mask, uint32_t epoch, uint32_t* d_error_word,
                     void* stream);

/* ------------------------------------------------------------------------------------------------
 * Sparse group: embedding tables sharded by rows (TF "div" partitioning), SURVEY.md 8(a) rows
 * a10..a12.  Reference: none -- TensorFlow's embedding_lookup over a partitioned variable and its
 * sparse optimizer ops on the PS tasks of k8s-operator.md:6.
 *
 * Shard s of the table lives on GPU s mod world.  d_shard_base[s] is row 0 of shard s; d_ids[w] /
 * d_grad_rows[w] are worker w's ids (int64) and gradient rows ([ids_per_worker, dim] fp32).  Any
 * of these may be peer-mapped pointers: the kernels read ids and rows where they live, so the
 * routing all-to-all is fused into them.
 *
 *   pull  out[k, :] = table[ids[k], :]                      (an out-of-range id gives a zero row)
 *   push  1. sort_segment: the entries (id, worker, position) of THIS GPU's shards, stably sorted
 *            by id -> sorted ids, source permutation, segment ids, unique ids, segment offsets
 *            (all in the workspace; counts stay on the device)
 *         2. apply: per unique id, G = (1/W) * fold of its gradient rows in sorted order (left
 *            fold; segments over 256 entries chunk-wise, see oracle/ps_oracle.c), then
 *            SGD     row -= lr * G
 *            Adagrad a += G*G; row -= (lr * G) / sqrt(a)       -- only touched rows change
 * Integer results are bit-exact against the oracle, and so are the rows (fixed fold order).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint64_t rows; /* rows of the whole table, < 2^32 */
    int dim;       /* floats per row: multiple of 4, 4..4096 */
    int nshards;   /* S, <= B200PS_MAX_WORKERS */
    int rank;      /* this GPU hosts the shards s with s % world == rank */
    int world;
} b200ps_table_geom_t;

int b200ps_sparse_gather(float* d_out, const float* const* d_shard_base, const b200ps_table_geom_t* geom,
                         const int64_t* d_ids, uint64_t n, void* stream);
/* Multi-GPU routing with posted NVLink writes (rows never travel as peer reads).
 * serve: the owner side of pull, for ALL workers at once -- this GPU walks every worker's ids,
 *   reads the rows of its own shards from local HBM and stores them into d_out[w] (worker w's
 *   [ids_per_worker, dim] buffer, a peer pointer when w lives elsewhere).  Every GPU of the box
 *   runs it; together they fill every d_out[w].  On one GPU it is the batched gather.
 * route: the worker side of push -- stores each gradient row whose id another GPU owns into
 *   d_dst_by_gpu[owner] + k*dim (that GPU's staging slot for this worker, k = position), so the
 *   owner's apply reads only local memory: it is given the staging slot as d_grad_rows[w] for
 *   every worker w that lives on another GPU. */
int b200ps_sparse_serve(float* const* d_out, const float* const* d_shard_base, const int64_t* const* d_ids,
                        int W, uint64_t ids_per_worker, const b200ps_table_geom_t* geom, void* stream);
int b200ps_sparse_route(const float* d_grad_rows, const int64_t* d_ids, uint64_t n,
                        const b200ps_table_geom_t* geom, float* const* d_dst_by_gpu, void* stream);
/* scratch for one push of n_total = W * ids_per_worker entries of a table with `dim` floats per
 * row (256-byte aligned device memory the caller owns; it holds the plan the apply call reads,
 * plus the per-chunk partial rows the apply call uses for ids with more than 256 entries).
 * sort_segment / apply take W * ids_per_worker < 2^32 - 2^16 entries (B200PS_ERR_RANGE above). */
size_t b200ps_sparse_workspace_bytes(uint64_t n_total, int dim);
int b200ps_sparse_sort_segment(const int64_t* const* d_ids, int W, uint64_t ids_per_worker,
                               const b200ps_table_geom_t* geom, void* d_workspace, size_t workspace_bytes,
                               void* stream);
int b200ps_sparse_apply_sgd(float* const* d_shard_base, const float* const* d_grad_rows, int W,
                            uint64_t ids_per_worker, const b200ps_table_geom_t* geom, float lr,
                            const void* d_workspace, void* stream);
int b200ps_sparse_apply_adagrad(float* const* d_shard_base, float* const* d_accum_base,
                                const float* const* d_grad_rows, int W, uint64_t ids_per_worker,
                                const b200ps_table_geom_t* geom, float lr, const void* d_workspace,
                                void* stream);
/* Debug accessor for the bit-exact checks: copies the plan of the last sort_segment to the host
 * (synchronises the stream).  Arrays may be NULL; sizes: n_kept for sorted ids / perm / segment
 * ids, n_unique for unique ids, n_unique + 1 for segment offsets (allocate n_total [+1]). */
int b200ps_sparse_plan_read(const void* d_workspace, uint64_t n_total, uint64_t rows, uint64_t* n_kept,
                            uint64_t* n_unique, uint64_t* n_bad, int64_t* h_sorted_ids, uint32_t* h_perm,
                            int32_t* h_seg_id, int64_t* h_unique_ids, uint32_t* h_seg_offset, void* stream);

/* No counterpart in the reference (nor in TensorFlow's PS ops): an entry point for callers that
 * must not allocate device memory.
 *
 * One whole sparse push with SGD -- exactly b200ps_sparse_sort_segment followed by
 * b200ps_sparse_apply_sgd -- whose workspace is a fixed array inside the library instead of
 * memory the caller provides.  It calls no cudaMalloc / cudaFree and no allocating library
 * routine; the caller passes only its inputs (ids, gradient rows) and the table shards it updates.
 * It is meant for a task whose HBM share was given out completely at launch (shards, gradient
 * buffers and replicas are sized to fill the GPU, so a later cudaMalloc can fail), and for a step
 * loop with push / apply / pull in flight on several streams, which cudaMalloc or cudaFree would
 * stall because they synchronise the whole device.
 *
 *  - limit: W * ids_per_worker <= B200PS_NOALLOC_MAX_ENTRIES, and the workspace of that push
 *    (b200ps_sparse_workspace_bytes, which grows with dim) must fit the array -- true for every
 *    dim <= 512 at the full entry count; anything larger is rejected with B200PS_ERR_RANGE (use
 *    the two-call form with your own workspace).
 *  - cost: every GPU this process uses the library on holds B200PS_NOALLOC_SCRATCH_BYTES of
 *    device memory for the array until the process exits.
 *  - the first call on each device looks up the array's device address (with that device
 *    current); make that first call before any stream capture begins.
 *  - the array is shared by all calls on a device and keeps the previous call's contents; the
 *    words that must start at zero are cleared on `stream` at the start of each call.  Two calls
 *    that can overlap on the GPU (different streams, no ordering between them) race on it:
 *    order them on one stream or with events.
 */
#define B200PS_NOALLOC_MAX_ENTRIES 262144ull
#define B200PS_NOALLOC_SCRATCH_BYTES 12582912ull /* 12 MiB: MAX_ENTRIES entries at dim <= 512 */
int b200ps_sparse_push_sgd_noalloc(float* const* d_shard_base, const int64_t* const* d_ids,
                                   const float* const* d_grad_rows, int W, uint64_t ids_per_worker,
                                   const b200ps_table_geom_t* geom, float lr, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Context group: the PS tasks and compute-task buffers of ONE GPU as one object the library owns
 * (SURVEY.md 8(b) "dense group").  Reference: none -- this is the part of TensorFlow's PS runtime
 * that holds the variables, their optimizer slots and the transport between the PS and WORKER
 * tasks of k8s-operator.md:6; a Go operator replacing pkg/trainer/replicas.go (images/tf3.PNG)
 * would bind exactly these calls through cgo.
 *
 * One process per GPU drives one context ("rank" r of "world" G, with the CUDA device it runs on
 * current on the calling thread).  Placement is b200ps_placement_gpu: shard s -> GPU s mod G,
 * compute task w -> GPU w mod G.  The context owns, in device memory:
 *   - one arena, exported to the other ranks of the box (CUDA IPC):
 *       [ replica: P_pad fp32 | (two-level, G > 1: this GPU's partial sum: P_pad) |
 *         one gradient buffer of P_pad per local compute task | barrier + slice-ready flags ]
 *   - the optimizer state (Adam m, v; momentum accumulator; FTRL accum, linear) of its shards,
 *   - the Adam step counter and the launch plan.
 * The master copy of slice s is the slice-s range of its owner's replica; the step kernel writes
 * the updated slice into every other rank's replica (pull fused into the step).
 *
 * Set-up sequence for G > 1 (the caller moves the 64-byte handles over any host channel and
 * provides ONE host-side barrier; the library calls nothing back):
 *     b200ps_ctx_create -> b200ps_ctx_export_handle -> (all-gather the handles) ->
 *     b200ps_ctx_import_handles -> (host barrier: every rank has imported) -> steps ...
 *     -> (host barrier: nobody is inside a step) -> b200ps_ctx_destroy
 * All ranks must issue the same sequence of collective calls (b200ps_dense_step,
 * b200ps_dense_step_async_ordered, b200ps_dense_step_host): each one contains stream-ordered
 * barriers through peer memory.  A rank that never arrives makes the others' barrier time out
 * (~4 s); from then on every kernel of the context is a no-op and every call returns
 * B200PS_ERR_STATE -- parameters are never updated from gradients that were not all there.
 *
 * Calls take the stream the work is ordered on.  A context is driven by one host thread.
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200ps_ctx b200ps_ctx_t;

enum { B200PS_OPT_SGD = 0, B200PS_OPT_ADAM = 1, B200PS_OPT_MOMENTUM = 2, B200PS_OPT_FTRL = 3,
       B200PS_OPT_ADAGRAD = 4 /* embedding tables only */ };
enum { B200PS_AGG_FLAT = 0,      /* g = (1/W) * left fold over the workers in worker order (a6) */
       B200PS_AGG_TWO_LEVEL = 1, /* G > 1: every GPU folds its own workers (worker order), the owner
                                    folds the per-GPU partial sums in GPU order; same bits as FLAT on one GPU */
       B200PS_AGG_MULTIMEM = 2   /* G > 1: as TWO_LEVEL, but the G partial sums are added inside the
                                    NVSwitch (multimem.ld_reduce) -- the switch's summation order, so
                                    results match the oracle to 1e-6 relative, not bit for bit.  Needs
                                    multicast support (b200ps_multicast_supported). */ };

typedef struct {
    int optimizer;       /* B200PS_OPT_* */
    int aggregate;       /* B200PS_AGG_* */
    double lr;           /* all optimizers */
    double beta1, beta2, eps; /* Adam (TF formulation) */
    double momentum;     /* ApplyMomentum */
    double ftrl_l1, ftrl_l2, ftrl_accum_init; /* ApplyFtrl with lr_power = -0.5 */
    int threads, ctas_per_sm, unroll; /* launch geometry of this context's dense kernels; 0 = default */
    int use_graph;       /* 1: b200ps_dense_step replays one CUDA graph per step (captured on first use) */
    int pipeline_folds;  /* G > 1, two-level: slice-wise fold that signals the slice owners, overlapping
                            the NVLink exchange with the local folds */
    uint64_t host_chunk_elems; /* pipelining granularity of b200ps_dense_step_host; 0 = 4 Mi elements */
} b200ps_dense_options_t;

/* fills *opt with the defaults: SGD, lr 0.01; Adam 1e-3 / 0.9 / 0.999 / 1e-8; momentum 0.9;
 * FTRL l1 = l2 = 0, accumulator 0.1; TWO_LEVEL aggregation; no graph. */
int b200ps_dense_options_default(b200ps_dense_options_t* opt);

/* S = b200ps_cluster_num_shards(cluster) must equal the map's shard count; W = the cluster's
 * compute tasks.  Allocates the arena and the state on the current device. */
int b200ps_ctx_create(const b200ps_cluster_t* cluster, const b200ps_shardmap_t* map, int rank, int world,
                      const b200ps_dense_options_t* opt, b200ps_ctx_t** out);
int b200ps_ctx_destroy(b200ps_ctx_t* ctx);
int b200ps_ctx_export_handle(b200ps_ctx_t* ctx, void* handle_out /* B200PS_IPC_HANDLE_BYTES */);
/* handles: world * B200PS_IPC_HANDLE_BYTES bytes, rank-major (this rank's own entry is ignored) */
int b200ps_ctx_import_handles(b200ps_ctx_t* ctx, const void* handles);
/* joins `stream` and the context's own copy streams, then reports a barrier time-out if one occurred */
int b200ps_ctx_sync(b200ps_ctx_t* ctx, void* stream);

/* The launch plan of one rank, pure host arithmetic (no device needed): what b200ps_ctx_create lays
 * out and what b200ps_dense_step executes. */
typedef struct b200ps_plan b200ps_plan_t;
int b200ps_plan_build(const b200ps_shardmap_t* map, int num_workers, int rank, int world, int aggregate,
                      b200ps_plan_t** out);
void b200ps_plan_destroy(b200ps_plan_t* plan);
uint64_t b200ps_plan_params_padded(const b200ps_plan_t* plan);          /* P_pad */
int b200ps_plan_two_level(const b200ps_plan_t* plan);                   /* 1 if per-GPU partial sums are used */
int b200ps_plan_first_grad_slot(const b200ps_plan_t* plan);             /* arena slot (units of P_pad) of the first gradient buffer */
uint64_t b200ps_plan_arena_elems(const b200ps_plan_t* plan, int of_rank); /* fp32 elements before the flag words */
int b200ps_plan_num_local_workers(const b200ps_plan_t* plan);
int b200ps_plan_local_worker(const b200ps_plan_t* plan, int i);         /* compute-task rank of the i-th local one */
int b200ps_plan_grad_offset(const b200ps_plan_t* plan, int worker, int* rank, uint64_t* offset);
int b200ps_plan_num_launches(const b200ps_plan_t* plan);                /* = local shards */
/* launch i: its shard and slice, the buffers it folds in fold order (rank, arena offset in elements;
 * arrays of B200PS_MAX_WORKERS) and the ranks whose replica it mirrors into (B200PS_MAX_MIRRORS).
 * Any output may be NULL. */
int b200ps_plan_launch(const b200ps_plan_t* plan, int i, int* shard, uint64_t* begin, uint64_t* end, int* n_src,
                       int* src_rank, uint64_t* src_offset, int* n_mirrors, int* mirror_rank);
const b200ps_plan_t* b200ps_ctx_plan(const b200ps_ctx_t* ctx);          /* owned by the context */
uint64_t b200ps_ctx_step_count(const b200ps_ctx_t* ctx);               /* optimizer steps taken (Adam's t) */
/* (4W + 8) n for SGD, (4W + 24) n for Adam and FTRL, (4W + 16) n momentum, summed over local slices */
uint64_t b200ps_ctx_algorithmic_bytes_per_step(const b200ps_ctx_t* ctx);

/* ---- parameters ---- */
int b200ps_dense_init(b200ps_ctx_t* ctx, const float* h_params /* P_pad */, void* stream); /* resets state and t */
int b200ps_dense_init_uniform(b200ps_ctx_t* ctx, uint64_t seed, void* stream);           /* p[i] = u(seed,0,0,i) */
/* device views (valid until destroy): the replica, a local compute task's gradient buffer (zero-copy
 * push: the worker's kernels write here), the state of a local shard (second is NULL if unused) */
int b200ps_dense_replica(b200ps_ctx_t* ctx, float** d_params);
int b200ps_dense_grad_buffer(b200ps_ctx_t* ctx, int worker, float** d_grad);
/* base of rank r's arena as mapped in this process (r = own rank: the local arena) -- for probes and
 * for workers that read a peer's replica or write a peer's gradient buffer directly */
int b200ps_ctx_arena_base(b200ps_ctx_t* ctx, int of_rank, float** d_base);
int b200ps_dense_state(b200ps_ctx_t* ctx, int shard, float** d_first, float** d_second, uint64_t* n);

/* ---- the step ---- */
/* push (H2D of one local compute task's gradient) / pull (D2H of the replica); asynchronous */
int b200ps_dense_push(b200ps_ctx_t* ctx, int worker, const float* h_grad, void* stream);
int b200ps_dense_pull(b200ps_ctx_t* ctx, int worker, float* h_params, void* stream);
/* push + aggregate + apply + pull for every local shard; gradients are already in the grad buffers */
int b200ps_dense_step(b200ps_ctx_t* ctx, void* stream);
/* a reproducible stand-in for asynchronous PS training: every worker's gradient
 * applied on its own, in worker order, no averaging -- W optimizer steps per call.  FLAT only. */
int b200ps_dense_step_async_ordered(b200ps_ctx_t* ctx, void* stream);
/* a whole step with HOST buffers, pipelined chunk by chunk over the context's copy streams:
 * h_grads[i] = gradient (P_pad floats, pinned) of the i-th local compute task; h_params_out (may be
 * NULL) receives the updated replica.  The copy-out of one step overlaps the copy-in of the next:
 * call b200ps_dense_host_sync before reading h_params_out. */
int b200ps_dense_step_host(b200ps_ctx_t* ctx, const float* const* h_grads, float* h_params_out, void* stream);
int b200ps_dense_host_sync(b200ps_ctx_t* ctx);

/* ---- inspection, checkpoint / resume ("storage" is a goal, k8s-operator.md:2) ---- */
/* h_params: the whole replica (P_pad); h_first / h_second (may be NULL): P_pad floats each, only the
 * ranges of this rank's shards are written.  Synchronises. */
int b200ps_dense_dump(b200ps_ctx_t* ctx, float* h_params, float* h_first, float* h_second, void* stream);
/* <directory>/ps-rank<r>-of-<G>.b200ps: this rank's shards (parameters, state), step counter and the
 * shard-map header; written to a temporary file and renamed.  Every rank calls save; put a host
 * barrier between the ranks' save and a later load (load reads every rank's file for the parameters). */
int b200ps_dense_save(b200ps_ctx_t* ctx, const char* directory, void* stream);
int b200ps_dense_load(b200ps_ctx_t* ctx, const char* directory, void* stream);

/* ---- tuning / tracing ---- */
int b200ps_ctx_set_tuning(b200ps_ctx_t* ctx, int threads, int ctas_per_sm, int unroll); /* 0 keeps a value */
/* switches graph replay of b200ps_dense_step on or off (as the use_graph option does at creation) */
int b200ps_ctx_set_graph(b200ps_ctx_t* ctx, int on);
/* profile on: b200ps_dense_step brackets its phases with events (and does not replay a graph);
 * b200ps_ctx_phase_ms returns the mean per step since the last call and resets (synchronises). */
int b200ps_ctx_set_profile(b200ps_ctx_t* ctx, int on);
int b200ps_ctx_phase_ms(b200ps_ctx_t* ctx, float* fold_ms, float* apply_ms, float* barrier_ms);
/* 1 if the current device and driver support NVSwitch multicast objects (B200PS_AGG_MULTIMEM) */
int b200ps_multicast_supported(void);

/* ------------------------------------------------------------------------------------------------
 * Table group: one embedding table of the job as an object the library owns (SURVEY.md 8(b) "sparse
 * group"; rows a10..a12).  Reference: none -- TensorFlow's partitioned embedding variable on the PS
 * tasks of k8s-operator.md:6, embedding_lookup and the sparse optimizer ops.
 *
 * Rows are partitioned by TF "div" over the cluster's PS shards; shard s lives on GPU s mod G,
 * compute task w on GPU w mod G.  A rank's table object owns ONE arena, exported to the other ranks:
 *   [ local shards | (Adagrad accumulators) | per local compute task: ids int64[B], gradient rows
 *     [B, dim], pulled rows [B, dim] | per REMOTE compute task: staging [B, dim] | barrier words ]
 * plus the push workspace.  Set-up for G > 1 is the context's: create -> export_handle -> (all-gather)
 * -> import_handles -> (host barrier) -> calls ... -> (host barrier) -> destroy.  Workers write ids
 * and gradient rows into their buffers (b200ps_table_worker_buffers) and read the pulled rows there.
 * Collective calls (every rank, same order): b200ps_sparse_pull, b200ps_sparse_push, b200ps_sparse_step.
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200ps_table b200ps_table_t;
typedef struct {
    int optimizer;     /* B200PS_OPT_SGD | B200PS_OPT_ADAGRAD */
    int use_graph;     /* 1: b200ps_sparse_step replays one CUDA graph (captured on first use) */
    double lr;
    double accum_init; /* Adagrad: TF's initial_accumulator_value, 0.1 */
} b200ps_table_options_t;
int b200ps_table_options_default(b200ps_table_options_t* opt);

/* The arena layout of rank `of_rank`, pure arithmetic (no device): byte offsets; entries that do not
 * live on that rank are UINT64_MAX.  Arrays: num_shards (shard, accum), num_workers (ids, grad,
 * pulled, staging -- the staging slot ON of_rank for a task placed elsewhere).  Any output may be NULL. */
int b200ps_table_layout(uint64_t rows, int dim, int num_shards, int num_workers, uint64_t ids_per_worker, int world,
                        int adagrad, int of_rank, uint64_t* arena_bytes, uint64_t* flag_offset, uint64_t* shard_offset,
                        uint64_t* accum_offset, uint64_t* ids_offset, uint64_t* grad_offset, uint64_t* pulled_offset,
                        uint64_t* staging_offset);

int b200ps_table_create(const b200ps_cluster_t* cluster, uint64_t rows, int dim, uint64_t ids_per_worker, int rank,
                        int world, const b200ps_table_options_t* opt, b200ps_table_t** out);
int b200ps_table_destroy(b200ps_table_t* table);
int b200ps_table_export_handle(b200ps_table_t* table, void* handle_out /* B200PS_IPC_HANDLE_BYTES */);
int b200ps_table_import_handles(b200ps_table_t* table, const void* handles /* world * B200PS_IPC_HANDLE_BYTES */);
int b200ps_table_sync(b200ps_table_t* table, void* stream); /* joins the stream, reports a barrier time-out */
int b200ps_table_init_uniform(b200ps_table_t* table, uint64_t seed, void* stream); /* row i, column d = u(seed,0,0,i*dim+d) */
/* device views: a local compute task's ids / gradient rows / pulled rows; a local shard's rows (and accumulators) */
int b200ps_table_worker_buffers(b200ps_table_t* table, int worker, int64_t** d_ids, float** d_grad_rows, float** d_pulled);
int b200ps_table_shard(b200ps_table_t* table, int shard, float** d_rows, float** d_accum, uint64_t* row_begin,
                       uint64_t* row_end);
/* a10 for every compute task of the job at once: afterwards the pulled buffer of every local task holds table[ids] */
int b200ps_sparse_pull(b200ps_table_t* table, void* stream);
/* a10 for one local task, rows read from the owning GPUs directly (not collective) */
int b200ps_sparse_pull_worker(b200ps_table_t* table, int worker, float* d_out, void* stream);
/* a11 + a12 for the shards of this GPU over every task's (ids, gradient rows) */
int b200ps_sparse_push(b200ps_table_t* table, void* stream);
/* pull of all tasks, then push, sharing one barrier */
int b200ps_sparse_step(b200ps_table_t* table, void* stream);
int b200ps_table_set_graph(b200ps_table_t* table, int on);
int b200ps_table_set_profile(b200ps_table_t* table, int on);
int b200ps_table_phase_ms(b200ps_table_t* table, float* serve_ms, float* route_sort_ms, float* apply_ms);
/* the plan of the last push (see b200ps_sparse_plan_read) */
int b200ps_table_plan_read(b200ps_table_t* table, uint64_t* n_kept, uint64_t* n_unique, uint64_t* n_bad,
                           int64_t* h_sorted_ids, uint32_t* h_perm, int32_t* h_seg_id, int64_t* h_unique_ids,
                           uint32_t* h_seg_offset, void* stream);
/* <directory>/table-rank<r>-of-<G>.b200ps: this rank's shards (and accumulators); temporary file + rename */
int b200ps_table_save(b200ps_table_t* table, const char* directory, void* stream);
int b200ps_table_load(b200ps_table_t* table, const char* directory, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200PS_H */
