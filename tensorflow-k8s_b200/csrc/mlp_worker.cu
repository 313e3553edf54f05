// mlp_worker.cu -- the compute tasks of BASELINE configs[1] on the GPU: forward and backward of the
// dense MLP (1024 -> 8192 -> 8192 -> 2048 -> 4096 by default) for every compute task placed on this
// GPU, writing each task's fp32 gradient straight into its gradient buffer of the PS arena
// (include/b200ps.h "MLP worker group").
//
// Reference: none -- "WORKER computes" (k8s-operator.md:6) is the user's TensorFlow program inside the
// WORKER containers.  This is the step on the other side of push / pull: with it a training step's host
// input is the mini-batch, not W gradient vectors, so nothing of size P crosses PCIe.
//
// Numerics: classic mixed precision.  Master parameters stay fp32 in the PS replica; once per step they
// are cast to bf16 (both W [in][out] and W^T [out][in], because every GEMM operand is fed K-major);
// activations are bf16; every contraction accumulates in fp32 (TMEM) on the 5th-gen tensor cores
// (gemm_sm100.cuh); gradients leave as fp32.  Loss = 1/(2B) sum ||y - t||^2 per compute task.
// Forward and the data-gradient GEMMs batch all local tasks (they share the weights); the weight
// gradient is one GEMM per layer whose tiles carry the task index (per-task K range and output).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <memory>
#include <vector>

#include "common.h"
#include "cuda_util.h"
#include "gemm_sm100.cuh"

namespace b200ps {
namespace {

using bf16 = __nv_bfloat16;

typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// the driver entry point is resolved through the runtime: the library does not link libcuda
encode_fn_t tensor_map_encoder() {
    static encode_fn_t fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
        return (encode_fn_t)p;
    }();
    return fn;
}

// a K-major bf16 matrix [rows][cols] (cols contiguous), boxes of 64 columns (128 bytes) x box_rows, 128-byte swizzle
int make_map(CUtensorMap* map, const bf16* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    encode_fn_t enc = tensor_map_encoder();
    B200PS_REQUIRE(enc, B200PS_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * sizeof(bf16)};
    const cuuint32_t box[2] = {(cuuint32_t)gemm::BK, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200PS_REQUIRE(r == CUDA_SUCCESS, B200PS_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d (matrix %llu x %llu)", (int)r,
                   (unsigned long long)rows, (unsigned long long)cols);
    return B200PS_OK;
}

template <int BN, int EPI>
int launch_gemm(const CUtensorMap& a, const CUtensorMap& b, const gemm::Args& g, int sms, cudaStream_t st) {
    using S = gemm::Smem<BN>;
    static bool configured = false;
    if (!configured) {
        B200PS_CUDA(cudaFuncSetAttribute(gemm::gemm_bf16_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
        configured = true;
    }
    const int tiles = (g.M / gemm::BM) * (g.N / BN) * g.batches;
    const int grid = tiles < sms ? tiles : sms;
    gemm::gemm_bf16_kernel<BN, EPI><<<grid, gemm::THREADS, S::TOTAL, st>>>(a, b, g);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

// fp32 W [rows][cols] -> bf16 W [rows][cols] and bf16 W^T [cols][rows]: 64 x 64 tiles through shared memory,
// 128-bit loads, 64-bit bf16x4 stores in both layouts (rows, cols multiples of 64)
__global__ void __launch_bounds__(256) cast_transpose_kernel(const float* __restrict__ w, bf16* __restrict__ wb, bf16* __restrict__ wbt,
                                                             int rows, int cols) {
    __shared__ float tile[64][65];
    const int c0 = blockIdx.x * 64, r0 = blockIdx.y * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 threads x 4 elements per row, 16 rows per pass
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int r = ty + 16 * p;
        const float4 x = __ldg(reinterpret_cast<const float4*>(w + (size_t)(r0 + r) * cols + c0) + tx);
        tile[r][4 * tx + 0] = x.x;
        tile[r][4 * tx + 1] = x.y;
        tile[r][4 * tx + 2] = x.z;
        tile[r][4 * tx + 3] = x.w;
        __nv_bfloat162 lo = __floats2bfloat162_rn(x.x, x.y), hi = __floats2bfloat162_rn(x.z, x.w);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&lo);
        o.y = *reinterpret_cast<uint32_t*>(&hi);
        *(reinterpret_cast<uint2*>(wb + (size_t)(r0 + r) * cols + c0) + tx) = o;
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int c = ty + 16 * p;  // output row of W^T
        __nv_bfloat162 lo = __floats2bfloat162_rn(tile[4 * tx + 0][c], tile[4 * tx + 1][c]);
        __nv_bfloat162 hi = __floats2bfloat162_rn(tile[4 * tx + 2][c], tile[4 * tx + 3][c]);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&lo);
        o.y = *reinterpret_cast<uint32_t*>(&hi);
        *(reinterpret_cast<uint2*>(wbt + (size_t)(c0 + c) * rows + r0) + tx) = o;
    }
}

// x [M][d] row-major -> xT [d][M]
__global__ void transpose_bf16_kernel(const bf16* __restrict__ x, bf16* __restrict__ xt, int rows, int cols) {
    __shared__ bf16 tile[32][34];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += 8) tile[i][threadIdx.x] = x[(size_t)(r0 + i) * cols + c0 + threadIdx.x];
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += 8) xt[(size_t)(c0 + i) * rows + r0 + threadIdx.x] = tile[threadIdx.x][i];
}

// dY = (y - t) / B (bf16, both layouts); per-block partial of sum (y - t)^2 in fixed order
__global__ void loss_grad_kernel(const bf16* __restrict__ y, const bf16* __restrict__ t, bf16* __restrict__ dy, bf16* __restrict__ dyT,
                                 float* __restrict__ partial, int M, int N, int batch) {
    // one block per (row tile of 32 rows, column tile of 32): grid (N/32, M/32), block (32, 8)
    __shared__ bf16 tile[32][34];
    __shared__ float red[8];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const float inv_b = 1.0f / (float)batch;
    float s = 0.f;
    for (int i = threadIdx.y; i < 32; i += 8) {
        const size_t at = (size_t)(r0 + i) * N + c0 + threadIdx.x;
        const float d = __bfloat162float(y[at]) - __bfloat162float(t[at]);
        s += d * d;
        const bf16 gq = __float2bfloat16_rn(d * inv_b);
        dy[at] = gq;
        tile[i][threadIdx.x] = gq;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) red[threadIdx.y] = s;
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += 8) dyT[(size_t)(c0 + i) * M + r0 + threadIdx.x] = tile[threadIdx.x][i];
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        float tot = 0.f;
        for (int k = 0; k < 8; ++k) tot += red[k];
        partial[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = tot;
    }
}

// loss[w] = 1/(2B) * sum of the partials of task w's row tiles, in index order (deterministic)
__global__ void loss_final_kernel(const float* __restrict__ partial, float* __restrict__ loss, int tiles_per_task, int batch) {
    const int w = blockIdx.x;
    __shared__ double acc[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < tiles_per_task; i += blockDim.x) s += (double)partial[(size_t)w * tiles_per_task + i];
    acc[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int k = 0; k < (int)blockDim.x; ++k) tot += acc[k];
        loss[w] = (float)(tot / (2.0 * (double)batch));
    }
}

// db[w][n] = sum over task w's rows of dY[m][n], read from dY^T [N][M] (rows contiguous): one warp per (n, w)
__global__ void bias_grad_kernel(const bf16* __restrict__ dyT, float* const* __restrict__ out, const uint64_t offset, int N, int M, int batch,
                                 int tasks) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= N * tasks) return;
    const int w = warp / N, n = warp - w * N;
    const bf16* row = dyT + (size_t)n * M + (size_t)w * batch;
    float s = 0.f;
    for (int i = lane; i < batch; i += 32) s += __bfloat162float(row[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[w][offset + n] = s;
}

}  // namespace
}  // namespace b200ps

using namespace b200ps;

struct b200ps_mlp {
    std::vector<int> dims;  // d_0 .. d_L
    int L = 0, batch = 0, tasks = 0, M = 0, sms = 0;
    std::vector<uint64_t> w_off, b_off;  // offsets of W_l, b_l in the flat parameter / gradient buffer
    // device buffers
    std::vector<bf16*> wb, wbt;          // W_l [d_l][d_l+1], W_l^T [d_l+1][d_l]
    std::vector<bf16*> act, actT;        // H_l [M][d_l], H_l^T [d_l][M]   (l = 0 .. L)
    std::vector<bf16*> dact, dactT;      // dH_l (l = 1 .. L)
    bf16* target = nullptr;
    float* partial = nullptr;
    float* loss = nullptr;
    float** d_grad_ptrs = nullptr;       // device copy of the per-task gradient pointers
    float* h_grad_ptrs[8] = {nullptr};   // what it holds
    // tensor maps: forward (A = H_l, B = W_l^T), dgrad (A = dH_l+1, B = W_l), wgrad (A = H_l^T, B = dH_l+1^T)
    std::vector<CUtensorMap> f_a, f_b, d_a, d_b, g_a, g_b;
    std::vector<int> f_bn, d_bn, g_bn;   // N tile per GEMM (256, or 128 when N is not a multiple of 256)
    std::vector<void*> owned;
    cudaStream_t s_copy = nullptr;       // the targets travel beside the forward pass
    cudaEvent_t ev_start = nullptr, ev_targets = nullptr;
};

namespace {

template <typename T>
int dev_alloc(b200ps_mlp* m, T** p, size_t n) {
    B200PS_CUDA(cudaMalloc(p, n * sizeof(T)));
    m->owned.push_back(*p);
    return B200PS_OK;
}

int pick_bn(int N) { return N % 256 == 0 ? 256 : 128; }

template <int EPI>
int run_gemm(int bn, const CUtensorMap& a, const CUtensorMap& b, const gemm::Args& g, int sms, cudaStream_t st) {
    return bn == 256 ? launch_gemm<256, EPI>(a, b, g, sms, st) : launch_gemm<128, EPI>(a, b, g, sms, st);
}

}  // namespace

extern "C" int b200ps_mlp_create(const int* dims, int n_dims, int batch_per_task, int local_tasks, const uint64_t* var_offsets,
                                 b200ps_mlp_t** out) {
    B200PS_REQUIRE(dims && out && var_offsets, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(n_dims >= 2 && n_dims <= 17, B200PS_ERR_RANGE, "an MLP needs 1..16 layers");
    B200PS_REQUIRE(local_tasks >= 1 && local_tasks <= 8, B200PS_ERR_RANGE, "1..8 compute tasks per GPU");
    B200PS_REQUIRE(batch_per_task >= 64 && batch_per_task % 64 == 0, B200PS_ERR_RANGE, "batch per task must be a multiple of 64");
    B200PS_REQUIRE((batch_per_task * local_tasks) % 128 == 0, B200PS_ERR_RANGE, "tasks x batch must be a multiple of 128 rows");
    for (int i = 0; i < n_dims; ++i)
        B200PS_REQUIRE(dims[i] >= 128 && dims[i] % 128 == 0, B200PS_ERR_RANGE, "layer width %d must be a multiple of 128", dims[i]);
    auto m = std::make_unique<b200ps_mlp>();
    m->dims.assign(dims, dims + n_dims);
    m->L = n_dims - 1;
    m->batch = batch_per_task;
    m->tasks = local_tasks;
    m->M = batch_per_task * local_tasks;
    B200PS_CUDA(sm_count(&m->sms));
    const int L = m->L, M = m->M;
    for (int l = 0; l < L; ++l) {
        m->w_off.push_back(var_offsets[2 * l]);
        m->b_off.push_back(var_offsets[2 * l + 1]);
        B200PS_REQUIRE(var_offsets[2 * l] % 4 == 0, B200PS_ERR_INVALID, "weight offset of layer %d is not 16-byte aligned", l);
    }
    m->wb.resize(L);
    m->wbt.resize(L);
    m->act.resize(L + 1);
    m->actT.resize(L + 1);
    m->dact.resize(L + 1, nullptr);
    m->dactT.resize(L + 1, nullptr);
    for (int l = 0; l < L; ++l) {
        const size_t n = (size_t)dims[l] * dims[l + 1];
        if (int rc = dev_alloc(m.get(), &m->wb[l], n)) return rc;
        if (int rc = dev_alloc(m.get(), &m->wbt[l], n)) return rc;
    }
    for (int l = 0; l <= L; ++l) {
        const size_t n = (size_t)M * dims[l];
        if (int rc = dev_alloc(m.get(), &m->act[l], n)) return rc;
        if (int rc = dev_alloc(m.get(), &m->actT[l], n)) return rc;
        if (l >= 1) {
            if (int rc = dev_alloc(m.get(), &m->dact[l], n)) return rc;
            if (int rc = dev_alloc(m.get(), &m->dactT[l], n)) return rc;
        }
    }
    if (int rc = dev_alloc(m.get(), &m->target, (size_t)M * dims[L])) return rc;
    if (int rc = dev_alloc(m.get(), &m->partial, (size_t)(M / 32) * (dims[L] / 32))) return rc;
    if (int rc = dev_alloc(m.get(), &m->loss, 8)) return rc;
    if (int rc = dev_alloc(m.get(), &m->d_grad_ptrs, 8)) return rc;
    m->f_a.resize(L);
    m->f_b.resize(L);
    m->d_a.resize(L);
    m->d_b.resize(L);
    m->g_a.resize(L);
    m->g_b.resize(L);
    m->f_bn.resize(L);
    m->d_bn.resize(L);
    m->g_bn.resize(L);
    for (int l = 0; l < L; ++l) {
        const int in = dims[l], on = dims[l + 1];
        // forward: H_{l+1}[M][on] = H_l[M][in] * (W_l^T [on][in])^T
        m->f_bn[l] = pick_bn(on);
        if (int rc = make_map(&m->f_a[l], m->act[l], M, in, gemm::BM)) return rc;
        if (int rc = make_map(&m->f_b[l], m->wbt[l], on, in, m->f_bn[l])) return rc;
        // dgrad: dH_l[M][in] = dH_{l+1}[M][on] * (W_l [in][on])^T
        m->d_bn[l] = pick_bn(in);
        if (l >= 1) {
            if (int rc = make_map(&m->d_a[l], m->dact[l + 1], M, on, gemm::BM)) return rc;
            if (int rc = make_map(&m->d_b[l], m->wb[l], in, on, m->d_bn[l])) return rc;
        }
        // wgrad (per task, K = its rows): dW_l[in][on] = H_l^T[in][M] * (dH_{l+1}^T [on][M])^T
        m->g_bn[l] = pick_bn(on);
        if (int rc = make_map(&m->g_a[l], m->actT[l], in, M, gemm::BM)) return rc;
        if (int rc = make_map(&m->g_b[l], m->dactT[l + 1], on, M, m->g_bn[l])) return rc;
    }
    B200PS_CUDA(cudaStreamCreateWithFlags(&m->s_copy, cudaStreamNonBlocking));
    B200PS_CUDA(cudaEventCreateWithFlags(&m->ev_start, cudaEventDisableTiming));
    B200PS_CUDA(cudaEventCreateWithFlags(&m->ev_targets, cudaEventDisableTiming));
    *out = m.release();
    return B200PS_OK;
}

extern "C" int b200ps_mlp_destroy(b200ps_mlp_t* m) {
    if (!m) return B200PS_OK;
    cudaDeviceSynchronize();
    if (m->s_copy) cudaStreamDestroy(m->s_copy);
    if (m->ev_start) cudaEventDestroy(m->ev_start);
    if (m->ev_targets) cudaEventDestroy(m->ev_targets);
    for (void* p : m->owned) cudaFree(p);
    delete m;
    return B200PS_OK;
}

extern "C" int b200ps_mlp_step(b200ps_mlp_t* m, const float* d_params, float* const* d_grads, const void* h_x, const void* h_t,
                               float* h_loss, void* stream) {
    B200PS_REQUIRE(m && d_params && d_grads && h_x && h_t, B200PS_ERR_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int L = m->L, M = m->M;
    const std::vector<int>& d = m->dims;
    // the step's host inputs: the mini-batches of the local tasks (features and targets), H2D
    // (the targets are not needed before the loss: they are copied on a side stream while the forward pass runs)
    B200PS_CUDA(cudaEventRecord(m->ev_start, st));  // behind the previous step's loss kernel, which read `target`
    B200PS_CUDA(cudaStreamWaitEvent(m->s_copy, m->ev_start, 0));
    B200PS_CUDA(cudaMemcpyAsync(m->target, h_t, (size_t)M * d[L] * sizeof(bf16), cudaMemcpyHostToDevice, m->s_copy));
    B200PS_CUDA(cudaEventRecord(m->ev_targets, m->s_copy));
    B200PS_CUDA(cudaMemcpyAsync(m->act[0], h_x, (size_t)M * d[0] * sizeof(bf16), cudaMemcpyHostToDevice, st));
    bool same = true;
    for (int w = 0; w < m->tasks; ++w) same = same && m->h_grad_ptrs[w] == d_grads[w];
    if (!same) {  // (a pageable source makes this copy synchronous: only when the buffers change)
        for (int w = 0; w < m->tasks; ++w) m->h_grad_ptrs[w] = d_grads[w];
        B200PS_CUDA(cudaMemcpyAsync(m->d_grad_ptrs, m->h_grad_ptrs, sizeof(float*) * (size_t)m->tasks, cudaMemcpyHostToDevice, st));
    }
    const dim3 tb(32, 8);
    transpose_bf16_kernel<<<dim3(d[0] / 32, M / 32), tb, 0, st>>>(m->act[0], m->actT[0], M, d[0]);
    // pull: the workers read the replica -- once per GPU, cast to the bf16 operand layouts
    for (int l = 0; l < L; ++l)
        cast_transpose_kernel<<<dim3(d[l + 1] / 64, d[l] / 64), 256, 0, st>>>(d_params + m->w_off[l], m->wb[l], m->wbt[l], d[l], d[l + 1]);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1 + (uint64_t)L, std::memory_order_relaxed);
    // forward
    for (int l = 0; l < L; ++l) {
        gemm::Args g{};
        g.M = M;
        g.N = d[l + 1];
        g.K = d[l];
        g.batches = 1;
        g.out = m->act[l + 1];
        g.outT = m->actT[l + 1];
        g.bias = d_params + m->b_off[l];
        g.relu = l + 1 < L ? 1 : 0;
        if (int rc = run_gemm<gemm::EPI_FWD>(m->f_bn[l], m->f_a[l], m->f_b[l], g, m->sms, st)) return rc;
    }
    // loss and its gradient
    B200PS_CUDA(cudaStreamWaitEvent(st, m->ev_targets, 0));
    loss_grad_kernel<<<dim3(d[L] / 32, M / 32), tb, 0, st>>>(m->act[L], m->target, m->dact[L], m->dactT[L], m->partial, M, d[L], m->batch);
    loss_final_kernel<<<m->tasks, 256, 0, st>>>(m->partial, m->loss, (m->batch / 32) * (d[L] / 32), m->batch);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(2, std::memory_order_relaxed);
    // backward
    for (int l = L - 1; l >= 0; --l) {
        {  // weight gradient of every task, straight into its gradient buffer
            gemm::Args g{};
            g.M = d[l];
            g.N = d[l + 1];
            g.K = m->batch;
            g.batches = m->tasks;
            g.k_batch_stride = m->batch;
            for (int w = 0; w < m->tasks; ++w) g.out32[w] = d_grads[w] + m->w_off[l];
            g.scale = 1.0f;  // dH already carries the 1/B of the mean
            if (int rc = run_gemm<gemm::EPI_WGRAD>(m->g_bn[l], m->g_a[l], m->g_b[l], g, m->sms, st)) return rc;
        }
        {
            const int warps = d[l + 1] * m->tasks;
            bias_grad_kernel<<<(warps * 32 + 255) / 256, 256, 0, st>>>(m->dactT[l + 1], m->d_grad_ptrs, m->b_off[l], d[l + 1], M, m->batch,
                                                                       m->tasks);
            B200PS_CUDA(cudaGetLastError());
            g_launches.fetch_add(1, std::memory_order_relaxed);
        }
        if (l >= 1) {  // data gradient, gated by the ReLU of layer l's output
            gemm::Args g{};
            g.M = M;
            g.N = d[l];
            g.K = d[l + 1];
            g.batches = 1;
            g.out = m->dact[l];
            g.outT = m->dactT[l];
            g.gate = m->act[l];
            if (int rc = run_gemm<gemm::EPI_DGRAD>(m->d_bn[l], m->d_a[l], m->d_b[l], g, m->sms, st)) return rc;
        }
    }
    if (h_loss) B200PS_CUDA(cudaMemcpyAsync(h_loss, m->loss, sizeof(float) * (size_t)m->tasks, cudaMemcpyDeviceToHost, st));
    return B200PS_OK;
}

// One stand-alone GEMM through the same kernel, for tests and roofline measurements:
// C[M][N] (bf16) = A[M][K] * B[N][K]^T, optional transposed copy; device pointers.
extern "C" int b200ps_gemm_bf16(const void* d_a, const void* d_b, void* d_c, void* d_ct, int M, int N, int K, void* stream) {
    B200PS_REQUIRE(d_a && d_b && d_c, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(M > 0 && N > 0 && K > 0 && M % 128 == 0 && N % 128 == 0 && K % 64 == 0, B200PS_ERR_RANGE,
                   "GEMM needs M, N multiples of 128 and K a multiple of 64 (got %d x %d x %d)", M, N, K);
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const int bn = pick_bn(N);
    CUtensorMap ma, mb;
    if (int rc = make_map(&ma, (const bf16*)d_a, (uint64_t)M, (uint64_t)K, gemm::BM)) return rc;
    if (int rc = make_map(&mb, (const bf16*)d_b, (uint64_t)N, (uint64_t)K, (uint32_t)bn)) return rc;
    gemm::Args g{};
    g.M = M;
    g.N = N;
    g.K = K;
    g.batches = 1;
    g.out = (bf16*)d_c;
    g.outT = (bf16*)d_ct;
    return run_gemm<gemm::EPI_FWD>(bn, ma, mb, g, sms, (cudaStream_t)stream);
}
