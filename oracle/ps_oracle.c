/*
 * ps_oracle.c -- CPU restatement of the parameter-server step.  TEST INFRASTRUCTURE ONLY.
 * PARITY UNPINNED (see ps_oracle.h): the reference holds no implementation of this path
 * (its only statement about it is k8s-operator.md:6); every function below follows the row of
 * SURVEY.md section 8(a) named in its comment.
 *
 * Build: gcc -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp (oracle/Makefile; v3 so the built .so runs on any host of the pool).
 * -ffp-contract=off is REQUIRED: one IEEE rounding per written operation.
 */
#include "ps_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

void ps_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n); /* launchers such as torchrun export OMP_NUM_THREADS=1 */
#else
    (void)n;
#endif
}

int ps_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------- synthetic inputs (8d) */

uint64_t ps_hash64(uint64_t seed, uint64_t step, uint64_t worker, uint64_t index) {
    /* counter packed by three odd multipliers, then the splitmix64 finaliser */
    uint64_t z = index ^ (seed * 0xD6E8FEB86659FD93ull) ^ (step * 0xA24BAED4963EE407ull) ^
                 (worker * 0x9FB21C651E98DF25ull);
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

float ps_uniform(uint64_t seed, uint64_t step, uint64_t worker, uint64_t index) {
    /* top 24 bits -> k * 2^-23 - 1, exact in binary32, range [-1, 1) */
    uint32_t k = (uint32_t)(ps_hash64(seed, step, worker, index) >> 40);
    return (float)k * 0x1p-23f - 1.0f;
}

void ps_fill_uniform(float* dst, uint64_t n, uint64_t seed, uint64_t step, uint64_t worker,
                     uint64_t first, int threads) {
    int64_t i;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (i = 0; i < (int64_t)n; ++i) dst[i] = ps_uniform(seed, step, worker, first + (uint64_t)i);
}

/* ---------------------------------------------------------------- a4: layout */

static uint64_t round_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

uint64_t ps_shard_map(const uint64_t* var_numel, int nvars, int nshards, uint32_t align,
                      uint64_t* var_offset, uint64_t* slice_begin, uint64_t* slice_end) {
    uint64_t off = 0, c;
    int v, s;
    for (v = 0; v < nvars; ++v) {
        if (var_offset) var_offset[v] = off;
        off += round_up(var_numel[v], align);
    }
    c = round_up((off + (uint64_t)nshards - 1) / (uint64_t)nshards, align);
    for (s = 0; s < nshards; ++s) {
        uint64_t b = (uint64_t)s * c, e = b + c;
        if (b > off) b = off;
        if (e > off) e = off;
        slice_begin[s] = b;
        slice_end[s] = e;
    }
    return off;
}

void ps_row_range(uint64_t rows, int nshards, int shard, uint64_t* begin, uint64_t* end) {
    uint64_t q = rows / (uint64_t)nshards, r = rows % (uint64_t)nshards, s = (uint64_t)shard;
    /* first r shards hold q+1 rows */
    *begin = s * q + (s < r ? s : r);
    *end = *begin + q + (s < r ? 1 : 0);
}

int ps_row_owner(uint64_t rows, int nshards, uint64_t id) {
    uint64_t q = rows / (uint64_t)nshards, r = rows % (uint64_t)nshards, t = r * (q + 1);
    if (id < t) return (int)(id / (q + 1));
    return (int)(r + (id - t) / q);
}

/* ---------------------------------------------------------------- a6..a8: dense */

#define BLK 512 /* elements per inner block; the per-element operation order never changes */

void ps_dense_sgd(float* p, const float* const* grads, int W, uint64_t n, float inv_w, float lr,
                  int threads) {
    int64_t nb = (int64_t)((n + BLK - 1) / BLK), b;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (b = 0; b < nb; ++b) {
        float acc[BLK];
        uint64_t i0 = (uint64_t)b * BLK, len = n - i0 < BLK ? n - i0 : BLK, i;
        int w;
        for (i = 0; i < len; ++i) acc[i] = grads[0][i0 + i];
        for (w = 1; w < W; ++w) {
            const float* g = grads[w] + i0;
            for (i = 0; i < len; ++i) acc[i] = acc[i] + g[i];
        }
        for (i = 0; i < len; ++i) {
            float g = acc[i] * inv_w;
            p[i0 + i] = p[i0 + i] - lr * g;
        }
    }
}

void ps_dense_momentum(float* p, float* accum, const float* const* grads, int W, uint64_t n, float inv_w, float lr,
                       float momentum, int threads) {
    int64_t nb = (int64_t)((n + BLK - 1) / BLK), b;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (b = 0; b < nb; ++b) {
        float acc[BLK];
        uint64_t i0 = (uint64_t)b * BLK, len = n - i0 < BLK ? n - i0 : BLK, i;
        int w;
        for (i = 0; i < len; ++i) acc[i] = grads[0][i0 + i];
        for (w = 1; w < W; ++w) {
            const float* g = grads[w] + i0;
            for (i = 0; i < len; ++i) acc[i] = acc[i] + g[i];
        }
        for (i = 0; i < len; ++i) {
            float g = acc[i] * inv_w;
            float an = momentum * accum[i0 + i] + g;
            accum[i0 + i] = an;
            p[i0 + i] = p[i0 + i] - lr * an;
        }
    }
}

/* TensorFlow's ApplyFtrl with lr_power = -0.5 (training_ops: FtrlCompute), one rounding per operation:
 *   na = accum + g*g;  sigma = (sqrt(na) - sqrt(accum)) / lr;  linear = linear + (g - sigma*p);
 *   x = sign(linear)*l1 - linear;  y = sqrt(na)/lr + 2*l2;  p = |linear| > l1 ? x/y : 0;  accum = na */
void ps_dense_ftrl(float* p, float* accum, float* linear, const float* const* grads, int W, uint64_t n, float inv_w,
                   float lr, float l1, float l2, int threads) {
    int64_t nb = (int64_t)((n + BLK - 1) / BLK), b;
    const float two_l2 = 2.0f * l2;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (b = 0; b < nb; ++b) {
        float acc[BLK];
        uint64_t i0 = (uint64_t)b * BLK, len = n - i0 < BLK ? n - i0 : BLK, i;
        int w;
        for (i = 0; i < len; ++i) acc[i] = grads[0][i0 + i];
        for (w = 1; w < W; ++w) {
            const float* g = grads[w] + i0;
            for (i = 0; i < len; ++i) acc[i] = acc[i] + g[i];
        }
        for (i = 0; i < len; ++i) {
            float g = acc[i] * inv_w;
            float a = accum[i0 + i], na = a + g * g;
            float sn = sqrtf(na);
            float sigma = (sn - sqrtf(a)) / lr;
            float sp = sigma * p[i0 + i];
            float li = linear[i0 + i] + (g - sp);
            float sgn = li > 0.f ? l1 : (li < 0.f ? -l1 : 0.f);
            float x = sgn - li;
            float y = sn / lr + two_l2;
            linear[i0 + i] = li;
            p[i0 + i] = fabsf(li) > l1 ? x / y : 0.f;
            accum[i0 + i] = na;
        }
    }
}

void ps_dense_fold(float* dst, const float* const* bufs, int W, uint64_t n, int threads) {
    /* the local half of the two-level aggregation: dst = ((b0 + b1) + b2) + ..., no scaling */
    int64_t nb = (int64_t)((n + BLK - 1) / BLK), b;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (b = 0; b < nb; ++b) {
        float acc[BLK];
        uint64_t i0 = (uint64_t)b * BLK, len = n - i0 < BLK ? n - i0 : BLK, i;
        int w;
        for (i = 0; i < len; ++i) acc[i] = bufs[0][i0 + i];
        for (w = 1; w < W; ++w) {
            const float* g = bufs[w] + i0;
            for (i = 0; i < len; ++i) acc[i] = acc[i] + g[i];
        }
        for (i = 0; i < len; ++i) dst[i0 + i] = acc[i];
    }
}

ps_adam_t ps_adam_scalars(double lr, double beta1, double beta2, double eps, uint64_t t) {
    ps_adam_t a;
    double b1t = pow(beta1, (double)t), b2t = pow(beta2, (double)t);
    a.lr_t = (float)(lr * sqrt(1.0 - b2t) / (1.0 - b1t));
    a.beta1 = (float)beta1;
    a.beta2 = (float)beta2;
    a.omb1 = (float)(1.0 - beta1);
    a.omb2 = (float)(1.0 - beta2);
    a.eps = (float)eps;
    return a;
}

void ps_dense_adam(float* p, float* m, float* v, const float* const* grads, int W, uint64_t n,
                   float inv_w, const ps_adam_t* a, int threads) {
    const float lr_t = a->lr_t, b1 = a->beta1, b2 = a->beta2, o1 = a->omb1, o2 = a->omb2,
                eps = a->eps;
    int64_t nb = (int64_t)((n + BLK - 1) / BLK), b;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (b = 0; b < nb; ++b) {
        float acc[BLK];
        uint64_t i0 = (uint64_t)b * BLK, len = n - i0 < BLK ? n - i0 : BLK, i;
        int w;
        for (i = 0; i < len; ++i) acc[i] = grads[0][i0 + i];
        for (w = 1; w < W; ++w) {
            const float* g = grads[w] + i0;
            for (i = 0; i < len; ++i) acc[i] = acc[i] + g[i];
        }
        for (i = 0; i < len; ++i) {
            float g = acc[i] * inv_w;
            float mn = b1 * m[i0 + i] + o1 * g;
            float vn = b2 * v[i0 + i] + o2 * (g * g);
            m[i0 + i] = mn;
            v[i0 + i] = vn;
            p[i0 + i] = p[i0 + i] - (lr_t * mn) / (sqrtf(vn) + eps);
        }
    }
}

/* ---------------------------------------------------------------- 8(f).1: linear-regression worker */

void ps_linreg_grad(float* grad, const float* params, uint64_t n, int batch, uint64_t seed_x, uint64_t seed_w,
                    uint64_t step, uint64_t worker, int threads) {
    const uint64_t nblk = (n + PS_LINREG_BLOCK - 1) / PS_LINREG_BLOCK;
    double* r = (double*)malloc((size_t)batch * sizeof(double));
    double* part = (double*)malloc((size_t)nblk * sizeof(double));
    const double inv_b = 1.0 / (double)batch;
    int b;
    int64_t i;
    for (b = 0; b < batch; ++b) {
        const uint64_t row = worker * (uint64_t)batch + (uint64_t)b;
        int64_t k;
        double acc = 0.0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
        for (k = 0; k < (int64_t)nblk; ++k) {
            uint64_t i0 = (uint64_t)k * PS_LINREG_BLOCK, i1 = i0 + PS_LINREG_BLOCK < n ? i0 + PS_LINREG_BLOCK : n, j;
            double a = 0.0;
            for (j = i0; j < i1; ++j) {
                /* (p - w_true) is formed in double: exact, both are floats */
                double d = (double)params[j] - (double)ps_uniform(seed_w, 0, 0, j);
                a = a + (double)ps_uniform(seed_x, step, row, j) * d;
            }
            part[k] = a;
        }
        for (k = 0; k < (int64_t)nblk; ++k) acc = acc + part[k];
        r[b] = acc;
    }
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(static)
    for (i = 0; i < (int64_t)n; ++i) {
        double acc = 0.0;
        int bb;
        for (bb = 0; bb < batch; ++bb)
            acc = acc + (double)ps_uniform(seed_x, step, worker * (uint64_t)batch + (uint64_t)bb, (uint64_t)i) * r[bb];
        grad[i] = (float)(inv_b * acc);
    }
    free(part);
    free(r);
}

/* ---------------------------------------------------------------- a10: gather */

void ps_gather(float* out, const float* table, const int64_t* ids, uint64_t n, int dim,
               int64_t row_base) {
    int64_t k;
#pragma omp parallel for schedule(static)
    for (k = 0; k < (int64_t)n; ++k)
        memcpy(out + k * (uint64_t)dim, table + (uint64_t)(ids[k] - row_base) * (uint64_t)dim,
               sizeof(float) * (size_t)dim);
}

/* ---------------------------------------------------------------- a11: bucket, sort, segment */

void ps_bucket_by_owner(const int64_t* ids, uint64_t n, uint64_t rows, int nshards,
                        uint32_t* perm, uint64_t* bucket_offset) {
    uint64_t k;
    int s;
    uint64_t* cursor = (uint64_t*)calloc((size_t)nshards + 1, sizeof(uint64_t));
    for (k = 0; k < n; ++k) cursor[ps_row_owner(rows, nshards, (uint64_t)ids[k]) + 1]++;
    for (s = 0; s < nshards; ++s) cursor[s + 1] += cursor[s];
    memcpy(bucket_offset, cursor, ((size_t)nshards + 1) * sizeof(uint64_t));
    for (k = 0; k < n; ++k) perm[cursor[ps_row_owner(rows, nshards, (uint64_t)ids[k])]++] = (uint32_t)k;
    free(cursor);
}

typedef struct {
    int64_t id;
    uint32_t src;
} keyed_t;

/* Stable LSD radix sort of (id, src) by id, 16 bits per pass (src ascends in the input, so a
 * stable sort by id alone gives the (id, worker, position) order). */
static void keyed_sort(keyed_t* a, uint64_t n) {
    keyed_t* b = (keyed_t*)malloc(n * sizeof(keyed_t));
    uint64_t* cnt = (uint64_t*)malloc(65537 * sizeof(uint64_t));
    uint64_t k, maxid = 0;
    int shift;
    for (k = 0; k < n; ++k)
        if ((uint64_t)a[k].id > maxid) maxid = (uint64_t)a[k].id;
    for (shift = 0; shift < 64 && (shift == 0 || (maxid >> shift) != 0); shift += 16) {
        keyed_t* t;
        memset(cnt, 0, 65537 * sizeof(uint64_t));
        for (k = 0; k < n; ++k) cnt[(((uint64_t)a[k].id >> shift) & 0xFFFF) + 1]++;
        for (k = 0; k < 65536; ++k) cnt[k + 1] += cnt[k];
        for (k = 0; k < n; ++k) b[cnt[((uint64_t)a[k].id >> shift) & 0xFFFF]++] = a[k];
        t = a; a = b; b = t;
        if (shift + 16 >= 64) { shift += 16; break; }
    }
    /* after an odd number of passes the result sits in the scratch buffer */
    if (((shift / 16) & 1) != 0) {
        memcpy(b, a, n * sizeof(keyed_t));
        free(a);
    } else {
        free(b);
    }
    free(cnt);
}

uint64_t ps_sort_segment(const int64_t* ids, uint64_t n, int64_t* sorted_ids, uint32_t* perm,
                         int32_t* seg_id, int64_t* unique_ids, uint32_t* seg_offset) {
    uint64_t k, nu = 0;
    keyed_t* a;
    if (n == 0) {
        if (seg_offset) seg_offset[0] = 0;
        return 0;
    }
    a = (keyed_t*)malloc(n * sizeof(keyed_t));
    for (k = 0; k < n; ++k) {
        a[k].id = ids[k];
        a[k].src = (uint32_t)k;
    }
    keyed_sort(a, n);
    for (k = 0; k < n; ++k) {
        if (k == 0 || a[k].id != a[k - 1].id) {
            unique_ids[nu] = a[k].id;
            seg_offset[nu] = (uint32_t)k;
            ++nu;
        }
        sorted_ids[k] = a[k].id;
        perm[k] = a[k].src;
        seg_id[k] = (int32_t)(nu - 1);
    }
    seg_offset[nu] = (uint32_t)n;
    free(a);
    return nu;
}

/* ---------------------------------------------------------------- a12: segmented apply */

/* Fixed-order segment sum of every column of one unique id, into total[dim].  Per column: segments
 * of at most PS_SEG_CHUNK entries are a plain left fold in sorted order; longer ones (hot ids under
 * Zipf skew) are folded chunk by chunk -- each chunk of PS_SEG_CHUNK consecutive entries left to
 * right, then the chunk sums left to right -- so that a GPU can give the chunks of one hot row to
 * different warps and still produce these bits.  (Rows are walked once, all columns together; the
 * columns are independent, so this is the same arithmetic as folding each column on its own.) */
static void segment_sum(float* total, float* acc, const float* grad_rows, const uint32_t* perm, uint32_t a, uint32_t b,
                        int dim) {
    uint32_t c0, j;
    int d;
    for (c0 = a; c0 < b; c0 += PS_SEG_CHUNK) {
        uint32_t c1 = b - c0 < PS_SEG_CHUNK ? b : c0 + PS_SEG_CHUNK;
        const float* r = grad_rows + (uint64_t)perm[c0] * (uint64_t)dim;
        for (d = 0; d < dim; ++d) acc[d] = r[d];
        for (j = c0 + 1; j < c1; ++j) {
            r = grad_rows + (uint64_t)perm[j] * (uint64_t)dim;
            for (d = 0; d < dim; ++d) acc[d] = acc[d] + r[d];
        }
        if (c0 == a)
            for (d = 0; d < dim; ++d) total[d] = acc[d];
        else
            for (d = 0; d < dim; ++d) total[d] = total[d] + acc[d];
    }
}

void ps_sparse_apply_sgd(float* table, int dim, int64_t row_base, const float* grad_rows,
                         const uint32_t* perm, const int64_t* unique_ids,
                         const uint32_t* seg_offset, uint64_t n_unique, float inv_w, float lr) {
    int64_t u;
    /* unique ids are distinct rows, so the loop over them is parallel; the fold inside one row is not */
#pragma omp parallel
    {
        float* total = (float*)malloc(2 * (size_t)dim * sizeof(float));
        float* acc = total + dim;
#pragma omp for schedule(dynamic, 64)
        for (u = 0; u < (int64_t)n_unique; ++u) {
            float* row = table + (uint64_t)(unique_ids[u] - row_base) * (uint64_t)dim;
            int d;
            segment_sum(total, acc, grad_rows, perm, seg_offset[u], seg_offset[u + 1], dim);
            for (d = 0; d < dim; ++d) {
                float g = total[d] * inv_w;
                row[d] = row[d] - lr * g;
            }
        }
        free(total);
    }
}

void ps_sparse_apply_adagrad(float* table, float* accum, int dim, int64_t row_base,
                             const float* grad_rows, const uint32_t* perm,
                             const int64_t* unique_ids, const uint32_t* seg_offset,
                             uint64_t n_unique, float inv_w, float lr) {
    int64_t u;
#pragma omp parallel
    {
        float* total = (float*)malloc(2 * (size_t)dim * sizeof(float));
        float* acc = total + dim;
#pragma omp for schedule(dynamic, 64)
        for (u = 0; u < (int64_t)n_unique; ++u) {
            uint64_t base = (uint64_t)(unique_ids[u] - row_base) * (uint64_t)dim;
            int d;
            segment_sum(total, acc, grad_rows, perm, seg_offset[u], seg_offset[u + 1], dim);
            for (d = 0; d < dim; ++d) {
                float g = total[d] * inv_w, an;
                an = accum[base + d] + g * g;
                accum[base + d] = an;
                table[base + d] = table[base + d] - (lr * g) / sqrtf(an);
            }
        }
        free(total);
    }
}
