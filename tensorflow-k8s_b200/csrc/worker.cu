// worker.cu -- a real worker for BASELINE configs[0]: the linear-regression gradient that feeds
// the dense PS path (SURVEY.md 8(f) rank 1 -- the step on the other side of push / pull).
//
// Reference: none.  In a TFJob this is the user's TensorFlow program inside the WORKER containers
// (k8s-operator.md:6: "WORKER computes"); it is outside the PS path proper, which is why it sits in
// a file of its own.  Features and targets are synthetic and generated in the kernels from the
// same integer hash as everything else, so no dataset is stored:
//   x_b[i] = u(seed_x, step, worker*B + b, i),  w_true[i] = u(seed_w, 0, 0, i),  y_b = x_b . w_true
//   r_b  = sum_i x_b[i] * (p[i] - w_true[i])      (double; 256-element blocks, then the block sums)
//   g[i] = (float)((1/B) * sum_b x_b[i] * r_b)    (double, b in order, one rounding to float)
// Explicit __dmul_rn / __dadd_rn: no FMA contraction, same bits as oracle/ps_oracle.c ps_linreg_grad.
#include <cuda_runtime.h>

#include "common.h"
#include "cuda_util.h"

namespace {

constexpr int BLOCK = 256;  // == PS_LINREG_BLOCK

__device__ __forceinline__ uint64_t mix(uint64_t key, uint64_t index) {
    uint64_t z = index ^ key;
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint64_t step, uint64_t worker) {
    return (seed * 0xD6E8FEB86659FD93ull) ^ (step * 0xA24BAED4963EE407ull) ^ (worker * 0x9FB21C651E98DF25ull);
}

__device__ __forceinline__ float uniform(uint64_t key, uint64_t index) {
    return __fsub_rn(__fmul_rn((float)(uint32_t)(mix(key, index) >> 40), 0x1p-23f), 1.0f);
}

// one thread per (row b, block k): left fold of 256 products in double
__global__ void linreg_block_kernel(const float* __restrict__ p, uint64_t n, int batch, uint64_t seed_x, uint64_t step,
                                    uint64_t row0, uint64_t key_w, uint64_t nblk, double* __restrict__ part) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nblk * (uint64_t)batch) return;
    const uint64_t b = t / nblk, k = t - b * nblk;
    const uint64_t key_x = stream_key(seed_x, step, row0 + b);
    const uint64_t i0 = k * BLOCK, i1 = i0 + BLOCK < n ? i0 + BLOCK : n;
    double a = 0.0;
    for (uint64_t j = i0; j < i1; ++j) {
        const double d = __dsub_rn((double)p[j], (double)uniform(key_w, j));
        a = __dadd_rn(a, __dmul_rn((double)uniform(key_x, j), d));
    }
    part[b * nblk + k] = a;
}

// one thread per row: the block sums, left to right
__global__ void linreg_residual_kernel(const double* __restrict__ part, uint64_t nblk, int batch, double* __restrict__ r) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= batch) return;
    double acc = 0.0;
    for (uint64_t k = 0; k < nblk; ++k) acc = __dadd_rn(acc, part[(uint64_t)b * nblk + k]);
    r[b] = acc;
}

// one thread per parameter: b = 0..B-1 in order
__global__ void linreg_grad_kernel(float* __restrict__ g, uint64_t n, int batch, uint64_t seed_x, uint64_t step, uint64_t row0,
                                   const double* __restrict__ r, double inv_b) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int b = 0; b < batch; ++b)
            acc = __dadd_rn(acc, __dmul_rn((double)uniform(stream_key(seed_x, step, row0 + (uint64_t)b), i), r[b]));
        g[i] = __double2float_rn(__dmul_rn(inv_b, acc));
    }
}

}  // namespace

extern "C" size_t b200ps_linreg_scratch_bytes(uint64_t n, int batch) {
    const uint64_t nblk = (n + BLOCK - 1) / BLOCK;
    return (size_t)((uint64_t)(batch > 0 ? batch : 0) * (nblk + 1) * sizeof(double));
}

extern "C" int b200ps_linreg_grad(float* d_grad, const float* d_params, uint64_t n, int batch, uint64_t seed_x,
                                  uint64_t seed_w, uint64_t step, uint64_t worker, void* d_scratch, size_t scratch_bytes,
                                  void* stream) {
    B200PS_REQUIRE(d_grad && d_params && d_scratch, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(n >= 1 && batch >= 1 && batch <= 65536, B200PS_ERR_RANGE, "n >= 1 and batch in 1..65536");
    B200PS_REQUIRE(((uintptr_t)d_scratch & 7) == 0, B200PS_ERR_INVALID, "scratch must be 8-byte aligned");
    B200PS_REQUIRE(scratch_bytes >= b200ps_linreg_scratch_bytes(n, batch), B200PS_ERR_RANGE,
                   "scratch of %zu bytes is too small: %zu needed", scratch_bytes, b200ps_linreg_scratch_bytes(n, batch));
    const uint64_t nblk = (n + BLOCK - 1) / BLOCK, row0 = worker * (uint64_t)batch;
    double* part = (double*)d_scratch;
    double* r = part + (uint64_t)batch * nblk;
    cudaStream_t st = (cudaStream_t)stream;
    const uint64_t work = nblk * (uint64_t)batch;
    linreg_block_kernel<<<(unsigned)((work + 127) / 128), 128, 0, st>>>(d_params, n, batch, seed_x, step, row0,
                                                                         stream_key(seed_w, 0, 0), nblk, part);
    linreg_residual_kernel<<<(batch + 63) / 64, 64, 0, st>>>(part, nblk, batch, r);
    int sms = 0;
    B200PS_CUDA(b200ps::sm_count(&sms));
    const uint64_t want = (n + 255) / 256, cap = (uint64_t)sms * 8;
    linreg_grad_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(d_grad, n, batch, seed_x, step, row0, r,
                                                                            1.0 / (double)batch);
    B200PS_CUDA(cudaGetLastError());
    b200ps::g_launches.fetch_add(3, std::memory_order_relaxed);
    return B200PS_OK;
}
