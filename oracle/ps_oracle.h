/*
 * ps_oracle.h -- CPU restatement of the parameter-server step.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED.  The reference (yph152/tensorflow-k8s) is documentation only; its single
 * statement about this path is the role model at k8s-operator.md:6 ("PS serves parameters, WORKER
 * computes").  It ships no implementation, no golden vector and no test for the path, so this file
 * restates the semantics fixed in SURVEY.md section 8(a) rows a4..a12 and is pinned only by the
 * hand-checkable known-answer tests of SURVEY.md section 8(c) (tests/test_oracle_kat.py) and by an
 * independent NumPy restatement (oracle/np_oracle.py).  It is a SUBSTITUTE CPU reference, not
 * the upstream project's code.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * link or call this.  The product (libb200ps.so) never does.
 *
 * All float arithmetic is IEEE-754 binary32, one rounding per written operation, no FMA
 * contraction (compile with -ffp-contract=off), so results are bit-comparable with the CUDA path.
 */
#ifndef PS_ORACLE_H
#define PS_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- synthetic inputs (SURVEY 8d "synthetic inputs"): integer hash -> fp32 in [-1, 1) ---- */
uint64_t ps_hash64(uint64_t seed, uint64_t step, uint64_t worker, uint64_t index);
float    ps_uniform(uint64_t seed, uint64_t step, uint64_t worker, uint64_t index);
/* dst[i] = ps_uniform(seed, step, worker, first + i) */
void ps_fill_uniform(float* dst, uint64_t n, uint64_t seed, uint64_t step, uint64_t worker,
                     uint64_t first, int threads);

/* ---- a4: layout ---- */
/* var_offset[v] = start of variable v in the flat buffer (each variable padded to `align`);
 * returns P_pad.  slice s = [slice_begin[s], slice_end[s]). */
uint64_t ps_shard_map(const uint64_t* var_numel, int nvars, int nshards, uint32_t align,
                      uint64_t* var_offset, uint64_t* slice_begin, uint64_t* slice_end);
/* TF "div" partitioning of `rows` rows over S shards. */
void    ps_row_range(uint64_t rows, int nshards, int shard, uint64_t* begin, uint64_t* end);
int     ps_row_owner(uint64_t rows, int nshards, uint64_t id);

/* ---- a6..a8: dense aggregate + apply, fixed worker order ---- */
/* g = inv_w * (((g0 + g1) + g2) + ...);  p -= lr * g */
void ps_dense_sgd(float* p, const float* const* grads, int W, uint64_t n, float inv_w, float lr,
                  int threads);
/* TF ApplyMomentum (no Nesterov): accum = momentum*accum + g;  p -= lr*accum */
void ps_dense_momentum(float* p, float* accum, const float* const* grads, int W, uint64_t n, float inv_w, float lr,
                       float momentum, int threads);
/* TF ApplyFtrl, lr_power = -0.5 (accum starts at 0.1, linear at 0); see ps_oracle.c for the operation order */
void ps_dense_ftrl(float* p, float* accum, float* linear, const float* const* grads, int W, uint64_t n, float inv_w,
                   float lr, float l1, float l2, int threads);
/* dst = ((b0 + b1) + b2) + ... (two-level aggregation: the per-GPU partial sum; the partials are
 * then folded in GPU order by ps_dense_* with inv_w = 1 / total workers) */
void ps_dense_fold(float* dst, const float* const* bufs, int W, uint64_t n, int threads);
typedef struct {
    float lr_t;   /* lr * sqrt(1 - beta2^t) / (1 - beta1^t), evaluated in double, rounded once */
    float beta1, beta2;
    float omb1, omb2; /* (float)(1.0 - (double)beta)  */
    float eps;
} ps_adam_t;
ps_adam_t ps_adam_scalars(double lr, double beta1, double beta2, double eps, uint64_t t);
/* m = b1*m + omb1*g;  v = b2*v + omb2*(g*g);  p -= (lr_t*m) / (sqrt(v) + eps) */
void ps_dense_adam(float* p, float* m, float* v, const float* const* grads, int W, uint64_t n,
                   float inv_w, const ps_adam_t* a, int threads);

/* ---- a10: sparse gather ---- */
void ps_gather(float* out, const float* table, const int64_t* ids, uint64_t n, int dim,
               int64_t row_base);

/* ---- a11: bucket + stable sort + segment ---- */
/* Stable bucket of ids by owner shard.  perm[j] = original position of the j-th routed id;
 * bucket_offset has nshards+1 entries. */
void ps_bucket_by_owner(const int64_t* ids, uint64_t n, uint64_t rows, int nshards,
                        uint32_t* perm, uint64_t* bucket_offset);
/* Stable sort of ids (input already in (worker, position) order).  Outputs: sorted_ids[n],
 * perm[n] (source position of sorted element j), seg_id[n], unique_ids[n_unique],
 * seg_offset[n_unique + 1].  Returns n_unique. */
uint64_t ps_sort_segment(const int64_t* ids, uint64_t n, int64_t* sorted_ids, uint32_t* perm,
                         int32_t* seg_id, int64_t* unique_ids, uint32_t* seg_offset);

/* ---- a12: segmented scatter-add + apply ---- */
#define PS_SEG_CHUNK 256u /* segments longer than this are folded chunk-wise (see ps_oracle.c) */
/* G_u = inv_w * fixed-order fold of grad rows in sorted order; SGD: row -= lr*G_u */
void ps_sparse_apply_sgd(float* table, int dim, int64_t row_base, const float* grad_rows,
                         const uint32_t* perm, const int64_t* unique_ids,
                         const uint32_t* seg_offset, uint64_t n_unique, float inv_w, float lr);
/* Adagrad: a += G_u*G_u;  row -= (lr*G_u) / sqrt(a) */
void ps_sparse_apply_adagrad(float* table, float* accum, int dim, int64_t row_base,
                             const float* grad_rows, const uint32_t* perm,
                             const int64_t* unique_ids, const uint32_t* seg_offset,
                             uint64_t n_unique, float inv_w, float lr);

/* ---- next row 8(f).1: a real worker for BASELINE configs[0] (linear regression) ----
 * Features x_b[i] = u(seed_x, step, worker*batch + b, i); targets y_b = x_b . w_true with
 * w_true[i] = u(seed_w, 0, 0, i).  Gradient of (1/2B) sum_b (x_b . p - y_b)^2:
 *   r_b  = sum_i x_b[i]*(p[i] - w_true[i])       in double, one rounding per multiply and per add:
 *                                                256-element blocks left to right, then the block sums
 *                                                left to right
 *   g[i] = (float)((1/B) * sum_b x_b[i]*r_b)     in double, b = 0..B-1 in order, one rounding to float
 * The worker is outside the PS path proper (last paragraph); it feeds it real gradients. */
#define PS_LINREG_BLOCK 256
void ps_linreg_grad(float* grad, const float* params, uint64_t n, int batch, uint64_t seed_x, uint64_t seed_w,
                    uint64_t step, uint64_t worker, int threads);
int ps_oracle_max_threads(void);
void ps_oracle_set_threads(int n); /* overrides OMP_NUM_THREADS for the calls that follow */

#ifdef __cplusplus
}
#endif
#endif
