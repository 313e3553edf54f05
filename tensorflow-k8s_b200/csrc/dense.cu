// dense.cu -- the dense PS step for one slice: push + aggregate + apply + pull in ONE kernel
// (include/b200ps.h "Kernel group").
//
// Reference: none -- the reference only names the roles (k8s-operator.md:6); in a TFJob the work
// is TensorFlow's gRPC tensor transport plus its ApplyGradientDescent / ApplyAdam ops on the PS
// task.  Here one sm_100a kernel streams W gradient buffers (local HBM or NVLink peer memory),
// folds them in fixed worker order, applies the optimizer to the slice and writes the new
// parameters to the master copy and to every mirror (the workers' copies on other GPUs).
//
// HBM-bound by construction (no contraction anywhere): per element the launch moves
//   SGD  4W + 8  bytes (W gradient reads, p read, p write)
//   Adam 4W + 24 bytes (+ m, v read and write)
// plus 4 bytes per mirror.  No tensor cores, no shared memory (nothing is reused): the design is
// 128-bit coalesced loads, all of a trip's loads issued before the first use, gradients marked
// read-once (L1 no-allocate, L2 evict-first) so they do not displace the resident state, and a
// grid that is a multiple of the SM count.
//
// Bit-exactness: every operation is an explicit round-to-nearest intrinsic (__fadd_rn, __fmul_rn,
// __fdiv_rn, __fsqrt_rn), which the compiler never contracts into FMA, in the operation order of
// oracle/ps_oracle.c.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdlib>
#include <map>
#include <mutex>

#include "common.h"
#include "cuda_util.h"
#include "dense_internal.h"

namespace b200ps {

std::atomic<uint64_t> g_launches{0};

namespace {

#define HAS_M(OPT) ((OPT) == OPT_ADAM || (OPT) == OPT_MOMENTUM || (OPT) == OPT_FTRL)  /* first state buffer: Adam's m / momentum accumulator / FTRL accum */
#define HAS_V(OPT) ((OPT) == OPT_ADAM || (OPT) == OPT_FTRL)  /* second: Adam's v / FTRL linear */

struct DenseParams {
    const float* g[B200PS_MAX_WORKERS];
    float* mirror[B200PS_MAX_MIRRORS];
    float* p;
    float* m;
    float* v;
    uint64_t n;   // elements in the slice
    uint64_t n4;  // float4 groups handled by the vector path (0 when a pointer is not 16-B aligned)
    int W;
    int n_mirror;
    float inv_w;
    float lr;  // SGD: lr.  Adam: lr_t.
    float beta1, beta2, omb1, omb2, eps;
    int load_kind;  // gradient loads: 0 = nc + L1 no-allocate + L2 evict-first, 1 = plain coherent ld.global
    int end_fence;  // 1: every thread ends with a system-scope fence (drains its peer stores inside the kernel)
    const uint32_t* abort_word;  // non-null and *abort_word != 0: a barrier of this context timed out -> do nothing
    // pipelined step: the sources of this slice are published by the producing ranks' fold kernels
    const uint32_t* wait_flags;  // non-null: local flag words, one per producing rank
    const uint32_t* wait_epoch;  // device word: the flags must reach *wait_epoch + 1
    uint32_t* abort_out;         // raised when the wait times out
    uint32_t wait_mask;          // producing ranks
};

__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// read-once stream (gradients): non-coherent path, no L1 allocation, first to leave L2
__device__ __forceinline__ float4 ld_once(const float4* ptr, uint64_t pol, int kind = 0) {
    float4 r;
    if (kind == 1) {
        asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(ptr));
        return r;
    }
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(ptr), "l"(pol));
    return r;
}

// state that this kernel rewrites (p, m, v): coherent load, no L1 allocation, normal L2 priority
__device__ __forceinline__ float4 ld_state(const float4* ptr) {
    float4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(ptr));
    return r;
}

__device__ __forceinline__ void st_f4(float4* ptr, const float4& v) {
    asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(ptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ float4 add4(const float4& a, const float4& b) {
    return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
}

// one element: g = inv_w * sum; optimizer update.  Same operation order as ps_dense_sgd / ps_dense_adam.
template <int OPT, typename PT>
__device__ __forceinline__ void apply1(float sum, float& p, float& m, float& v, const PT& a, const float lr) {
    if (OPT == OPT_FOLD) {
        p = sum;
        return;
    }
    const float g = __fmul_rn(sum, a.inv_w);
    if (OPT == OPT_SGD) {
        p = __fsub_rn(p, __fmul_rn(lr, g));
    } else if (OPT == OPT_MOMENTUM) {  // TF ApplyMomentum: accum = momentum*accum + g; p -= lr*accum
        m = __fadd_rn(__fmul_rn(a.beta1, m), g);
        p = __fsub_rn(p, __fmul_rn(lr, m));
    } else if (OPT == OPT_FTRL) {
        // TF ApplyFtrl, lr_power = -0.5 (m = accum, v = linear; beta1 = l1, beta2 = 2*l2); order as ps_dense_ftrl
        const float na = __fadd_rn(m, __fmul_rn(g, g));
        const float sn = __fsqrt_rn(na);
        const float sigma = __fdiv_rn(__fsub_rn(sn, __fsqrt_rn(m)), lr);
        v = __fadd_rn(v, __fsub_rn(g, __fmul_rn(sigma, p)));
        const float x = __fsub_rn(v > 0.f ? a.beta1 : (v < 0.f ? -a.beta1 : 0.f), v);
        const float y = __fadd_rn(__fdiv_rn(sn, lr), a.beta2);
        p = fabsf(v) > a.beta1 ? __fdiv_rn(x, y) : 0.f;
        m = na;
    } else {
        m = __fadd_rn(__fmul_rn(a.beta1, m), __fmul_rn(a.omb1, g));
        v = __fadd_rn(__fmul_rn(a.beta2, v), __fmul_rn(a.omb2, __fmul_rn(g, g)));
        p = __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, m), __fadd_rn(__fsqrt_rn(v), a.eps)));
    }
}

template <int OPT, typename PT>
__device__ __forceinline__ void apply4(const float4& s, float4& p, float4& m, float4& v, const PT& a, const float lr) {
    apply1<OPT>(s.x, p.x, m.x, v.x, a, lr);
    apply1<OPT>(s.y, p.y, m.y, v.y, a, lr);
    apply1<OPT>(s.z, p.z, m.z, v.z, a, lr);
    apply1<OPT>(s.w, p.w, m.w, v.w, a, lr);
}

// One trip: U float4 groups per stream at i, i+stride, ..., all loads issued before any use.
// WT > 0: W known at compile time (whole fold unrolled, W*U gradient loads in flight).
// WT == 0: runtime W, folded four buffers at a time.
template <int OPT, int WT, int U>
__device__ __forceinline__ void trip(const DenseParams& a, const float lr, uint64_t i, uint64_t stride, uint64_t pol) {
    float4 acc[U], P[U], M[U], V[U];
    if (WT > 0) {
        float4 G[WT > 0 ? WT : 1][U];
#pragma unroll
        for (int w = 0; w < WT; ++w)
#pragma unroll
            for (int u = 0; u < U; ++u) G[w][u] = ld_once(reinterpret_cast<const float4*>(a.g[w]) + i + u * stride, pol, a.load_kind);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (OPT != OPT_FOLD) P[u] = ld_state(reinterpret_cast<const float4*>(a.p) + i + u * stride);
            if (HAS_M(OPT)) M[u] = ld_state(reinterpret_cast<const float4*>(a.m) + i + u * stride);
            if (HAS_V(OPT)) V[u] = ld_state(reinterpret_cast<const float4*>(a.v) + i + u * stride);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            acc[u] = G[0][u];
#pragma unroll
            for (int w = 1; w < WT; ++w) acc[u] = add4(acc[u], G[w][u]);
        }
    } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (OPT != OPT_FOLD) P[u] = ld_state(reinterpret_cast<const float4*>(a.p) + i + u * stride);
            if (HAS_M(OPT)) M[u] = ld_state(reinterpret_cast<const float4*>(a.m) + i + u * stride);
            if (HAS_V(OPT)) V[u] = ld_state(reinterpret_cast<const float4*>(a.v) + i + u * stride);
            acc[u] = ld_once(reinterpret_cast<const float4*>(a.g[0]) + i + u * stride, pol, a.load_kind);
        }
        for (int w0 = 1; w0 < a.W; w0 += 4) {
            float4 G[4][U];
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (w0 + c < a.W) {
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        G[c][u] = ld_once(reinterpret_cast<const float4*>(a.g[w0 + c]) + i + u * stride, pol, a.load_kind);
                }
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (w0 + c < a.W) {
#pragma unroll
                    for (int u = 0; u < U; ++u) acc[u] = add4(acc[u], G[c][u]);
                }
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        apply4<OPT>(acc[u], P[u], M[u], V[u], a, lr);
        const uint64_t j = i + u * stride;
        st_f4(reinterpret_cast<float4*>(a.p) + j, P[u]);
        if (HAS_M(OPT)) st_f4(reinterpret_cast<float4*>(a.m) + j, M[u]);
        if (HAS_V(OPT)) st_f4(reinterpret_cast<float4*>(a.v) + j, V[u]);
        for (int k = 0; k < a.n_mirror; ++k) st_f4(reinterpret_cast<float4*>(a.mirror[k]) + j, P[u]);
    }
}

template <int OPT, int WT, int U>
__global__ void dense_apply_kernel(const __grid_constant__ DenseParams a) {
    if (a.abort_word && *(const volatile uint32_t*)a.abort_word) return;  // never update from gradients that did not all arrive
    if (a.wait_flags) {  // pipelined step: wait until every producing rank has published this slice's partial sum
        __shared__ uint32_t dead;
        if (threadIdx.x == 0) {
            const uint32_t want = *(const volatile uint32_t*)a.wait_epoch + 1u;
            uint32_t bad = 0;
            for (int r = 0; r < 32 && !bad; ++r) {
                if (!((a.wait_mask >> r) & 1u)) continue;
                const volatile uint32_t* f = a.wait_flags + r;
                const long long t0 = clock64();
                while ((int32_t)(*f - want) < 0) {
                    if (clock64() - t0 > 8000000000ll || *(const volatile uint32_t*)a.abort_word) {  // ~4 s: a producer died
                        *(volatile uint32_t*)a.abort_out = 1u;
                        bad = 1;
                        break;
                    }
                }
            }
            __threadfence_system();  // acquire: the partial sums behind the flags
            dead = bad;
        }
        __syncthreads();
        if (dead) return;
    }
    // the step size stays a kernel parameter (constant bank): reading it from device memory into a
    // register cost 4 % of the C1 step on B200 (0.619 -> 0.648 ms); graph replay re-arms the node instead
    const float lr = a.lr;
    const uint64_t pol = policy_evict_first();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t i = tid;
    for (; i + (U - 1) * stride < a.n4; i += U * stride) trip<OPT, WT, U>(a, lr, i, stride, pol);
    if (U > 1)
        for (; i < a.n4; i += stride) trip<OPT, WT, 1>(a, lr, i, stride, pol);
    // scalar tail: n % 4 elements, or the whole slice when a pointer is not 16-byte aligned
    for (uint64_t j = a.n4 * 4 + tid; j < a.n; j += stride) {
        float s = a.g[0][j];
        for (int w = 1; w < a.W; ++w) s = __fadd_rn(s, a.g[w][j]);
        float p = OPT == OPT_FOLD ? 0.f : a.p[j], m = 0.f, v = 0.f;
        if (HAS_M(OPT)) m = a.m[j];
        if (HAS_V(OPT)) v = a.v[j];
        apply1<OPT>(s, p, m, v, a, lr);
        a.p[j] = p;
        if (HAS_M(OPT)) a.m[j] = m;
        if (HAS_V(OPT)) a.v[j] = v;
        for (int k = 0; k < a.n_mirror; ++k) a.mirror[k][j] = p;
    }
    if (a.end_fence) __threadfence_system();
}

__device__ __forceinline__ uint64_t hash64(uint64_t key, uint64_t index) {
    uint64_t z = index ^ key;
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void fill_uniform_kernel(float* __restrict__ dst, uint64_t n, uint64_t key, uint64_t first) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t k = (uint32_t)(hash64(key, first + i) >> 40);
        dst[i] = __fsub_rn(__fmul_rn((float)k, 0x1p-23f), 1.0f);
    }
}

// ------------------------------------------------------------------------------------------------
// B200PS_AGG_MULTIMEM: the G partial sums of a slice are added INSIDE the NVSwitch -- one
// multimem.ld_reduce on the multicast mapping returns their sum, so 4 bytes per element arrive at the
// owner instead of 4 (G - 1) -- and the updated slice reaches every replica (this GPU's included) with
// one multimem.st.  The summation order is the switch's: checked against the oracle at 1e-6 relative.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 mc_ld_reduce(const float4* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void mc_st(float4* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int OPT>
__global__ void __launch_bounds__(256) dense_apply_mc_kernel(const __grid_constant__ McApply a) {
    if (*(const volatile uint32_t*)a.abort_word) return;
    if (a.wait_flags) {  // pipelined step: every rank has published its partial sum of this unit
        __shared__ uint32_t dead;
        if (threadIdx.x == 0) {
            const uint32_t want = *(const volatile uint32_t*)a.wait_epoch + 1u;
            uint32_t bad = 0;
            for (int r = 0; r < 32 && !bad; ++r) {
                if (!((a.wait_mask >> r) & 1u)) continue;
                const volatile uint32_t* f = a.wait_flags + r;
                const long long t0 = clock64();
                while ((int32_t)(*f - want) < 0) {
                    if (clock64() - t0 > 8000000000ll || *(const volatile uint32_t*)a.abort_word) {
                        *(volatile uint32_t*)a.abort_out = 1u;
                        bad = 1;
                        break;
                    }
                }
            }
            __threadfence_system();
            dead = bad;
        }
        __syncthreads();
        if (dead) return;
    }
    const float lr = a.lr;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, n4 = a.n / 4;
    for (uint64_t i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += 2 * stride) {
        const bool two = i0 + stride < n4;
        // the ranks walk their slices from different relative offsets (rot4), wrapping around
        uint64_t i = i0 + a.rot4, j = i0 + stride + a.rot4;
        if (i >= n4) i -= n4;
        if (j >= n4) j -= n4;
        float4 s0 = mc_ld_reduce(reinterpret_cast<const float4*>(a.g_mc) + i), s1 = s0;
        if (two) s1 = mc_ld_reduce(reinterpret_cast<const float4*>(a.g_mc) + j);
        float4 P0 = ld_state(reinterpret_cast<const float4*>(a.p) + i), P1 = P0, M0 = P0, M1 = P0, V0 = P0, V1 = P0;
        if (HAS_M(OPT)) M0 = ld_state(reinterpret_cast<const float4*>(a.m) + i);
        if (HAS_V(OPT)) V0 = ld_state(reinterpret_cast<const float4*>(a.v) + i);
        if (two) {
            P1 = ld_state(reinterpret_cast<const float4*>(a.p) + j);
            if (HAS_M(OPT)) M1 = ld_state(reinterpret_cast<const float4*>(a.m) + j);
            if (HAS_V(OPT)) V1 = ld_state(reinterpret_cast<const float4*>(a.v) + j);
        }
        apply4<OPT>(s0, P0, M0, V0, a, lr);
        mc_st(reinterpret_cast<float4*>(a.p_mc) + i, P0);
        if (HAS_M(OPT)) st_f4(reinterpret_cast<float4*>(a.m) + i, M0);
        if (HAS_V(OPT)) st_f4(reinterpret_cast<float4*>(a.v) + i, V0);
        if (two) {
            apply4<OPT>(s1, P1, M1, V1, a, lr);
            mc_st(reinterpret_cast<float4*>(a.p_mc) + j, P1);
            if (HAS_M(OPT)) st_f4(reinterpret_cast<float4*>(a.m) + j, M1);
            if (HAS_V(OPT)) st_f4(reinterpret_cast<float4*>(a.v) + j, V1);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Pipelined two-level aggregation: ONE fold kernel walks all S slices in order; when the last CTA has
// finished slice s it publishes this rank's partial sum of s to the slice's owner (a flag word in the
// owner's arena, a posted NVLink store), and the owner's apply kernel -- already resident, waiting in
// its prologue -- starts on s while this kernel is folding s + 1.  HBM-bound fold and NVLink-bound
// exchange overlap without extra launches.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dense_fold_publish_kernel(const __grid_constant__ FoldPublish f) {
    if (*(const volatile uint32_t*)f.abort_word) return;
    const uint64_t pol = policy_evict_first();
    const uint32_t epoch = *(volatile uint32_t*)f.epoch + 1u;  // every CTA reads it before the last arrival of the last slice
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int s = 0; s < f.n_slices; ++s) {
        const uint64_t b4 = f.begin[s] / 4, e4 = f.end[s] / 4;  // slices are 1024-element aligned
        for (uint64_t i = b4 + tid; i < e4; i += 2 * stride) {
            const bool two = i + stride < e4;
            float4 a0 = ld_once(reinterpret_cast<const float4*>(f.g[0]) + i, pol), a1 = a0;
            if (two) a1 = ld_once(reinterpret_cast<const float4*>(f.g[0]) + i + stride, pol);
            for (int w0 = 1; w0 < f.W; w0 += 4) {
                float4 x[4], y[4];
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (w0 + c < f.W) {
                        x[c] = ld_once(reinterpret_cast<const float4*>(f.g[w0 + c]) + i, pol);
                        if (two) y[c] = ld_once(reinterpret_cast<const float4*>(f.g[w0 + c]) + i + stride, pol);
                    }
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (w0 + c < f.W) {
                        a0 = add4(a0, x[c]);
                        if (two) a1 = add4(a1, y[c]);
                    }
            }
            st_f4(reinterpret_cast<float4*>(f.out) + i, a0);
            if (two) st_f4(reinterpret_cast<float4*>(f.out) + i + stride, a1);
        }
        __threadfence_system();  // this thread's part of the partial sum is visible to the peers
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint32_t arrived = atomicAdd(&f.done[s], 1u);
            if (arrived == gridDim.x - 1) {  // last CTA of slice s
                f.done[s] = 0;               // re-armed for the next step
                __threadfence_system();
                if (f.flag[s]) *(volatile uint32_t*)f.flag[s] = epoch;
                if (s == f.n_slices - 1) *(volatile uint32_t*)f.epoch = epoch;  // every CTA has read the old value by now
            }
        }
    }
}

Tuning g_tuning;  // process-wide default of the stateless entry points; a context carries its own

// largest CTA the kernel's register count allows (1024 threads x > 64 registers does not fit an SM)
template <typename Kern>
int max_threads_for(Kern kern) {
    cudaFuncAttributes fa{};
    return cudaFuncGetAttributes(&fa, kern) == cudaSuccess ? fa.maxThreadsPerBlock : 256;
}

// kernel function -> OPT_*, filled as instantiations are first launched: lets a context find the
// apply nodes of a captured graph (dense_node_opt) and re-arm their step size (dense_node_set_lr)
constexpr int MC_NODE = 0x100;  // the node's argument is a McApply, not a DenseParams
std::mutex g_registry_mutex;
std::map<const void*, int> g_registry;

template <int OPT, int WT, int U>
cudaError_t launch_k(const DenseParams& a, int grid, int threads, cudaStream_t st) {
    static const int limit = max_threads_for(dense_apply_kernel<OPT, WT, U>);
    static const bool registered = [] {
        std::lock_guard<std::mutex> lock(g_registry_mutex);
        g_registry[(const void*)dense_apply_kernel<OPT, WT, U>] = OPT;
        return true;
    }();
    (void)registered;
    if (threads > limit) {  // keep the thread count of the launch constant: more, smaller CTAs
        const int t2 = limit / 32 * 32;
        grid = (int)(((long long)grid * threads + t2 - 1) / t2);
        threads = t2;
    }
    dense_apply_kernel<OPT, WT, U><<<grid, threads, 0, st>>>(a);
    return cudaGetLastError();
}

template <int OPT, int WT>
cudaError_t launch_u(const DenseParams& a, int grid, int threads, int unroll, cudaStream_t st) {
    switch (unroll) {
        case 1: return launch_k<OPT, WT, 1>(a, grid, threads, st);
        case 4: return launch_k<OPT, WT, 4>(a, grid, threads, st);
        default: return launch_k<OPT, WT, 2>(a, grid, threads, st);
    }
}

template <int OPT>
cudaError_t launch_w(const DenseParams& a, int grid, int threads, int unroll, cudaStream_t st) {
    switch (a.W) {
        case 1: return launch_u<OPT, 1>(a, grid, threads, unroll, st);
        case 2: return launch_u<OPT, 2>(a, grid, threads, unroll, st);
        case 4: return launch_u<OPT, 4>(a, grid, threads, unroll, st);
        case 8: return launch_u<OPT, 8>(a, grid, threads, unroll > 2 ? 2 : unroll, st);
        default: return launch_u<OPT, 0>(a, grid, threads, unroll > 2 ? 2 : unroll, st);
    }
}

int dense_apply(int opt, DenseParams& a, const float* const* d_grads, float* const* d_mirrors, void* stream,
                const Tuning& t = g_tuning) {
    B200PS_REQUIRE(a.p && d_grads, B200PS_ERR_INVALID, "null parameter or gradient pointer");
    B200PS_REQUIRE(a.W >= 1 && a.W <= B200PS_MAX_WORKERS, B200PS_ERR_RANGE, "W = %d is outside 1..%d", a.W,
                   B200PS_MAX_WORKERS);
    B200PS_REQUIRE(a.n_mirror >= 0 && a.n_mirror <= B200PS_MAX_MIRRORS, B200PS_ERR_RANGE,
                   "n_mirrors = %d is outside 0..%d", a.n_mirror, B200PS_MAX_MIRRORS);
    B200PS_REQUIRE(a.n_mirror == 0 || d_mirrors, B200PS_ERR_INVALID, "null mirror list");
    if (opt == OPT_ADAM) B200PS_REQUIRE(a.m && a.v, B200PS_ERR_INVALID, "Adam needs m and v");
    if (opt == OPT_MOMENTUM) B200PS_REQUIRE(a.m, B200PS_ERR_INVALID, "momentum needs the accumulator");
    if (opt == OPT_FTRL) B200PS_REQUIRE(a.m && a.v, B200PS_ERR_INVALID, "FTRL needs accum and linear");
    if (a.n == 0) return B200PS_OK;

    uintptr_t bits = (uintptr_t)a.p | (uintptr_t)a.m | (uintptr_t)a.v;
    for (int w = 0; w < a.W; ++w) {
        B200PS_REQUIRE(d_grads[w], B200PS_ERR_INVALID, "gradient buffer %d is null", w);
        a.g[w] = d_grads[w];
        bits |= (uintptr_t)d_grads[w];
    }
    for (int k = 0; k < a.n_mirror; ++k) {
        B200PS_REQUIRE(d_mirrors[k], B200PS_ERR_INVALID, "mirror %d is null", k);
        a.mirror[k] = d_mirrors[k];
        bits |= (uintptr_t)d_mirrors[k];
    }
    a.n4 = (bits & 15) ? 0 : a.n / 4;

    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    static const int env_load_kind = getenv("B200PS_DENSE_LOAD") ? atoi(getenv("B200PS_DENSE_LOAD")) : 0;  // experiments
    if (env_load_kind) a.load_kind = env_load_kind;
    static const int env_end_fence = getenv("B200PS_DENSE_END_FENCE") ? atoi(getenv("B200PS_DENSE_END_FENCE")) : 0;  // experiments
    a.end_fence = (env_end_fence && a.n_mirror > 0) ? 1 : 0;
    const uint64_t work = a.n4 ? a.n4 : a.n;
    uint64_t want = (work + (uint64_t)t.threads * (uint64_t)t.unroll - 1) / ((uint64_t)t.threads * (uint64_t)t.unroll);
    const uint64_t cap = (uint64_t)sms * (uint64_t)t.ctas_per_sm;
    const int grid = (int)(want < cap ? (want ? want : 1) : cap);
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = opt == OPT_SGD    ? launch_w<OPT_SGD>(a, grid, t.threads, t.unroll, st)
                    : opt == OPT_ADAM ? launch_w<OPT_ADAM>(a, grid, t.threads, t.unroll, st)
                    : opt == OPT_MOMENTUM ? launch_w<OPT_MOMENTUM>(a, grid, t.threads, t.unroll, st)
                    : opt == OPT_FTRL ? launch_w<OPT_FTRL>(a, grid, t.threads, t.unroll, st)
                                      : launch_w<OPT_FOLD>(a, grid, t.threads, t.unroll, st);
    B200PS_CUDA(e);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

}  // namespace

int dense_launch(const DenseLaunch& l, const Tuning& t, cudaStream_t stream) {
    DenseParams a{};
    a.p = l.p;
    a.m = l.s1;
    a.v = l.s2;
    a.n = l.n;
    a.W = l.W;
    a.inv_w = l.inv_w;
    a.n_mirror = l.n_mirrors;
    a.lr = l.lr;
    a.beta1 = l.c1;
    a.beta2 = l.c2;
    a.omb1 = l.omb1;
    a.omb2 = l.omb2;
    a.eps = l.eps;
    a.abort_word = l.abort_word;
    a.wait_flags = l.wait_flags;
    a.wait_epoch = l.wait_epoch;
    a.wait_mask = l.wait_mask;
    a.abort_out = l.abort_out;
    a.load_kind = l.wait_flags ? 1 : 0;  // sources published during this kernel's lifetime are read with coherent loads
    return dense_apply(l.opt, a, l.grads, l.mirrors, (void*)stream, t);
}

int dense_launch_mc(const McApply& a, int ctas_per_sm, cudaStream_t stream) {
    B200PS_REQUIRE(a.g_mc && a.p_mc && a.p && a.abort_word && a.n % 4 == 0 &&
                       (((uintptr_t)a.g_mc | (uintptr_t)a.p_mc | (uintptr_t)a.p | (uintptr_t)a.m | (uintptr_t)a.v) & 15) == 0,
                   B200PS_ERR_INVALID, "bad multicast-apply arguments");
    if (a.n == 0) return B200PS_OK;
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const uint64_t want = (a.n / 4 + 511) / 512, cap = (uint64_t)sms * (uint64_t)ctas_per_sm;
    const int grid = (int)(want < cap ? want : cap);
    cudaStream_t st = stream;
    static const bool registered = [] {  // graph replay re-arms the step size of these nodes too (dense_node_set_lr)
        std::lock_guard<std::mutex> lock(g_registry_mutex);
        g_registry[(const void*)dense_apply_mc_kernel<OPT_ADAM>] = OPT_ADAM | MC_NODE;
        return true;
    }();
    (void)registered;
    switch (a.opt) {
        case OPT_ADAM: dense_apply_mc_kernel<OPT_ADAM><<<grid, 256, 0, st>>>(a); break;
        case OPT_MOMENTUM: dense_apply_mc_kernel<OPT_MOMENTUM><<<grid, 256, 0, st>>>(a); break;
        case OPT_FTRL: dense_apply_mc_kernel<OPT_FTRL><<<grid, 256, 0, st>>>(a); break;
        default: dense_apply_mc_kernel<OPT_SGD><<<grid, 256, 0, st>>>(a);
    }
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

int dense_fold_publish(const FoldPublish& f, int ctas_per_sm, cudaStream_t stream) {
    B200PS_REQUIRE(f.W >= 1 && f.W <= B200PS_MAX_WORKERS && f.n_slices >= 1 && f.n_slices <= 64 && f.out && f.done && f.epoch &&
                       f.abort_word,
                   B200PS_ERR_INVALID, "bad fold-publish arguments");
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    dense_fold_publish_kernel<<<sms * ctas_per_sm, 256, 0, stream>>>(f);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

int dense_node_opt(cudaGraphNode_t node) {
    cudaGraphNodeType ty;
    if (cudaGraphNodeGetType(node, &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) return -1;
    cudaKernelNodeParams kp{};
    if (cudaGraphKernelNodeGetParams(node, &kp) != cudaSuccess) return -1;
    std::lock_guard<std::mutex> lock(g_registry_mutex);
    auto it = g_registry.find(kp.func);
    return it == g_registry.end() ? -1 : (it->second & 0xff);
}

cudaError_t dense_node_set_lr(cudaGraphExec_t exec, cudaGraphNode_t node, float lr) {
    cudaKernelNodeParams kp{};
    cudaError_t e = cudaGraphKernelNodeGetParams(node, &kp);
    if (e != cudaSuccess) return e;
    bool mc = false;
    {
        std::lock_guard<std::mutex> lock(g_registry_mutex);
        auto it = g_registry.find(kp.func);
        mc = it != g_registry.end() && (it->second & MC_NODE);
    }
    DenseParams a{};
    McApply b{};
    void* args[1];
    if (mc) {
        b = *static_cast<const McApply*>(kp.kernelParams[0]);  // the node's copy of the launch arguments
        b.lr = lr;
        args[0] = &b;
    } else {
        a = *static_cast<const DenseParams*>(kp.kernelParams[0]);
        a.lr = lr;
        args[0] = &a;
    }
    kp.kernelParams = args;
    return cudaGraphExecKernelNodeSetParams(exec, node, &kp);
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int b200ps_adam_scalars(double lr, double beta1, double beta2, double eps, uint64_t t, b200ps_adam_t* out) {
    B200PS_REQUIRE(out, B200PS_ERR_INVALID, "null argument");
    B200PS_REQUIRE(t >= 1, B200PS_ERR_RANGE, "Adam step t starts at 1");
    const double b1t = std::pow(beta1, (double)t), b2t = std::pow(beta2, (double)t);
    out->lr_t = (float)(lr * std::sqrt(1.0 - b2t) / (1.0 - b1t));
    out->beta1 = (float)beta1;
    out->beta2 = (float)beta2;
    out->omb1 = (float)(1.0 - beta1);
    out->omb2 = (float)(1.0 - beta2);
    out->eps = (float)eps;
    return B200PS_OK;
}

extern "C" int b200ps_dense_fold(float* d_out, const float* const* d_bufs, int W, uint64_t n, void* stream) {
    DenseParams a{};
    a.p = d_out;
    a.n = n;
    a.W = W;
    return dense_apply(OPT_FOLD, a, d_bufs, nullptr, stream);
}

extern "C" int b200ps_dense_apply_sgd(float* d_params, const float* const* d_grads, int W, uint64_t n, float inv_w,
                                      float lr, float* const* d_mirrors, int n_mirrors, void* stream) {
    DenseParams a{};
    a.p = d_params;
    a.n = n;
    a.W = W;
    a.inv_w = inv_w;
    a.n_mirror = n_mirrors;
    a.lr = lr;
    return dense_apply(OPT_SGD, a, d_grads, d_mirrors, stream);
}

extern "C" int b200ps_dense_apply_adam(float* d_params, float* d_m, float* d_v, const float* const* d_grads, int W,
                                       uint64_t n, float inv_w, const b200ps_adam_t* opt, float* const* d_mirrors,
                                       int n_mirrors, void* stream) {
    B200PS_REQUIRE(opt, B200PS_ERR_INVALID, "null optimizer scalars");
    DenseParams a{};
    a.p = d_params;
    a.m = d_m;
    a.v = d_v;
    a.n = n;
    a.W = W;
    a.inv_w = inv_w;
    a.n_mirror = n_mirrors;
    a.lr = opt->lr_t;
    a.beta1 = opt->beta1;
    a.beta2 = opt->beta2;
    a.omb1 = opt->omb1;
    a.omb2 = opt->omb2;
    a.eps = opt->eps;
    return dense_apply(OPT_ADAM, a, d_grads, d_mirrors, stream);
}

extern "C" int b200ps_dense_apply_momentum(float* d_params, float* d_accum, const float* const* d_grads, int W, uint64_t n,
                                           float inv_w, float lr, float momentum, float* const* d_mirrors, int n_mirrors,
                                           void* stream) {
    DenseParams a{};
    a.p = d_params;
    a.m = d_accum;
    a.n = n;
    a.W = W;
    a.inv_w = inv_w;
    a.n_mirror = n_mirrors;
    a.lr = lr;
    a.beta1 = momentum;
    return dense_apply(OPT_MOMENTUM, a, d_grads, d_mirrors, stream);
}

extern "C" int b200ps_dense_apply_ftrl(float* d_params, float* d_accum, float* d_linear, const float* const* d_grads, int W,
                                       uint64_t n, float inv_w, float lr, float l1, float l2, float* const* d_mirrors,
                                       int n_mirrors, void* stream) {
    B200PS_REQUIRE(lr > 0.f && l1 >= 0.f && l2 >= 0.f, B200PS_ERR_RANGE, "FTRL needs lr > 0, l1 >= 0, l2 >= 0");
    DenseParams a{};
    a.p = d_params;
    a.m = d_accum;
    a.v = d_linear;
    a.n = n;
    a.W = W;
    a.inv_w = inv_w;
    a.n_mirror = n_mirrors;
    a.lr = lr;
    a.beta1 = l1;
    a.beta2 = 2.0f * l2;
    return dense_apply(OPT_FTRL, a, d_grads, d_mirrors, stream);
}

extern "C" int b200ps_fill_uniform(float* d_dst, uint64_t n, uint64_t seed, uint64_t step, uint64_t worker,
                                   uint64_t first, void* stream) {
    B200PS_REQUIRE(d_dst || n == 0, B200PS_ERR_INVALID, "null destination");
    if (n == 0) return B200PS_OK;
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const uint64_t key = (seed * 0xD6E8FEB86659FD93ull) ^ (step * 0xA24BAED4963EE407ull) ^ (worker * 0x9FB21C651E98DF25ull);
    const uint64_t want = (n + 255) / 256, cap = (uint64_t)sms * 8;
    fill_uniform_kernel<<<(int)(want < cap ? want : cap), 256, 0, (cudaStream_t)stream>>>(d_dst, n, key, first);
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_set_dense_tuning(int threads, int ctas_per_sm, int unroll) {
    if (threads) {
        B200PS_REQUIRE(threads >= 32 && threads <= 1024 && threads % 32 == 0, B200PS_ERR_RANGE,
                       "threads must be a multiple of 32 in 32..1024");
        g_tuning.threads = threads;
    }
    if (ctas_per_sm) {
        B200PS_REQUIRE(ctas_per_sm >= 1 && ctas_per_sm <= 32, B200PS_ERR_RANGE, "ctas_per_sm must be in 1..32");
        g_tuning.ctas_per_sm = ctas_per_sm;
    }
    if (unroll) {
        B200PS_REQUIRE(unroll == 1 || unroll == 2 || unroll == 4, B200PS_ERR_RANGE, "unroll must be 1, 2 or 4");
        g_tuning.unroll = unroll;
    }
    return B200PS_OK;
}

extern "C" uint64_t b200ps_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
