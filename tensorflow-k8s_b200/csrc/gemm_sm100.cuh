// gemm_sm100.cuh -- a hand-written sm_100a GEMM for the MLP worker: C[M,N] = A[M,K] * B[N,K]^T, bf16
// operands (both K-major), fp32 accumulation in tensor memory.
//
// Reference: none -- the WORKER side of k8s-operator.md:6 ("WORKER computes") is the user's
// TensorFlow program; this is the contraction of BASELINE configs[1]'s dense MLP.
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0      TMA producer: cp.async.bulk.tensor loads of the A (128 x 64) and B (BN x 64) tiles into a
//               ring of shared-memory stages (128-byte swizzle), completion on mbarriers
//   warp 1      MMA issuer: one lane issues tcgen05.mma (cta_group::1, kind::f16, M = 128, N = BN, K = 16)
//               from shared-memory descriptors into one of TWO accumulator buffers in TMEM;
//               tcgen05.commit releases the stage / publishes the accumulator
//   warps 2..5  epilogue: tcgen05.ld (32 lanes x 32 columns per warp and instruction) -> registers ->
//               fused epilogue -> global memory; overlaps the next tile's main loop (TMEM double buffer)
// Tiles are walked n-major so that CTAs running at the same time share the B (weight) tile in L2.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace b200ps {
namespace gemm {

constexpr int BM = 128;       // UMMA M
constexpr int BK = 64;        // bf16 elements per stage row = 128 bytes = one swizzle atom
constexpr int UMMA_K = 16;    // fixed for 16-bit operands
constexpr int THREADS = 192;  // 6 warps

enum Epilogue {
    EPI_FWD = 0,    // out = bf16(act(acc + bias[n])) to out[M][N] and outT[N][M]; act = ReLU when relu != 0
    EPI_DGRAD = 1,  // out = bf16(acc * (gate[m][n] > 0)) to out[M][N] and outT[N][M]
    EPI_WGRAD = 2   // out32[z][M][N] = acc * scale   (z = compute task: per-task K range and output pointer)
};

struct Args {
    int M, N, K;            // problem; M % 128 == 0, N % BN == 0, K % 64 == 0
    int batches;            // EPI_WGRAD: number of compute tasks (K ranges); else 1
    int k_batch_stride;     // EPI_WGRAD: K offset (elements) between tasks in A and B
    __nv_bfloat16* out;     // [M][N]
    __nv_bfloat16* outT;    // [N][M] (may be null)
    const float* bias;      // [N] (EPI_FWD, may be null)
    const __nv_bfloat16* gate;  // [M][N] activations of the layer (EPI_DGRAD, may be null = no gate)
    float* out32[8];        // EPI_WGRAD: one output per task
    float scale;            // EPI_WGRAD
    int relu;               // EPI_FWD
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c_inner, int c_outer) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
        : "memory");
}

// shared-memory matrix descriptor of a K-major tile stored with the 128-byte swizzle: rows of 128 bytes,
// 8-row groups 1024 bytes apart (SBO), leading-dimension offset unused (1), descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                  // LBO (16-byte units)
    d |= (uint64_t)(1024 >> 4) << 32;        // SBO
    d |= (uint64_t)1 << 46;                  // version
    d |= (uint64_t)2 << 61;                  // SWIZZLE_128B
    return d;
}

// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// completion of all previously issued MMAs arrives on the mbarrier (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

template <int BN>
struct Smem {
    static constexpr int STAGES = BN == 256 ? 4 : 6;
    static constexpr int A_BYTES = BM * BK * 2;  // 16 KB
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
    static constexpr int STAGING_OFFSET = BAR_OFFSET + 256;          // EPI_WGRAD: 4 warps x 32 rows x 36 floats
    static constexpr int STAGING_BYTES = 4 * 32 * 36 * 4;
    static constexpr int TOTAL = STAGING_OFFSET + STAGING_BYTES + 1024;  // + slack for 1024-byte alignment
};

template <int BN, int EPI>
__global__ void __launch_bounds__(THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const __grid_constant__ Args g) {
    using S = Smem<BN>;
    constexpr int STAGES = S::STAGES;
    constexpr uint32_t TMEM_COLS = 2 * BN;  // two accumulator buffers of BN fp32 columns (512 or 256: powers of two)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);  // swizzle-128B tiles need 1024-byte alignment
    uint64_t* full = (uint64_t*)(smem + S::BAR_OFFSET);
    uint64_t* empty = full + STAGES;
    uint64_t* acc_full = empty + STAGES;
    uint64_t* acc_empty = acc_full + 2;
    uint32_t* tmem_ptr = (uint32_t*)(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = g.M / BM, n_tiles = g.N / BN;
    const int tiles_per_batch = m_tiles * n_tiles;
    const int total_tiles = tiles_per_batch * g.batches;
    const int k_blocks = g.K / BK;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tma_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tma_b) : "memory");
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 4);  // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // the allocating warp also frees
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int z = tile / tiles_per_batch, t = tile - z * tiles_per_batch;
                const int n_blk = t / m_tiles, m_blk = t - n_blk * m_tiles;
                const int k0 = z * g.k_batch_stride;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);  // the MMAs that read this slot have completed
                    uint8_t* a_dst = smem + stage * S::STAGE_BYTES;
                    mbar_expect_tx(&full[stage], S::STAGE_BYTES);
                    tma_load_2d(a_dst, &tma_a, &full[stage], k0 + kb * BK, m_blk * BM);
                    tma_load_2d(a_dst + S::A_BYTES, &tma_b, &full[stage], k0 + kb * BK, n_blk * BN);
                    if (++stage == STAGES) stage = 0, phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
                const int acc = it & 1;
                mbar_wait(&acc_empty[acc], ((it >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&full[stage], phase);  // TMA has landed this stage
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(smem + stage * S::STAGE_BYTES);
                    const uint64_t adesc = umma_desc_k_sw128(a_addr), bdesc = umma_desc_k_sw128(a_addr + S::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)  // 32 bytes further along K inside the swizzle atom: +2 in the address field
                        umma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (uint32_t)(kb | k));
                    umma_commit(&empty[stage]);  // frees the slot when those MMAs are done
                    if (++stage == STAGES) stage = 0, phase ^= 1;
                }
                umma_commit(&acc_full[acc]);  // accumulator complete
            }
        }
    } else {
        // ===== epilogue: warps 2..5 own TMEM lanes 32*(warp % 4) .. +31 =====
        const int lane_grp = warp & 3;
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
            const int z = tile / tiles_per_batch, t = tile - z * tiles_per_batch;
            const int n_blk = t / m_tiles, m_blk = t - n_blk * m_tiles;
            const int acc = it & 1;
            mbar_wait(&acc_full[acc], (it >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int m = m_blk * BM + lane_grp * 32 + lane;  // the accumulator row this thread holds
            const uint32_t t_row = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld32(t_row + (uint32_t)(c * 32), v);
                const int n0 = n_blk * BN + c * 32;
                if (EPI == EPI_WGRAD) {
                    // fp32 rows straight into the task's gradient buffer.  A thread holds one ROW of the tile, so
                    // storing from registers would touch 32 different rows per instruction (16-byte pieces); the
                    // 32 x 32 block is turned through a per-warp shared-memory tile instead (conflict-free both
                    // ways with a 36-float pitch) and leaves as full 128-byte lines, four rows per instruction.
                    float* stg = reinterpret_cast<float*>(smem + S::STAGING_OFFSET) + lane_grp * (32 * 36);
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(stg + lane * 36 + j) =
                            make_float4(__uint_as_float(v[j]) * g.scale, __uint_as_float(v[j + 1]) * g.scale,
                                        __uint_as_float(v[j + 2]) * g.scale, __uint_as_float(v[j + 3]) * g.scale);
                    __syncwarp();
                    float* dst = g.out32[z] + (size_t)(m_blk * BM + lane_grp * 32) * g.N + n0;
                    const int sub = lane >> 3, col = (lane & 7) * 4;
#pragma unroll
                    for (int r4 = 0; r4 < 32; r4 += 4) {
                        const float4 o = *reinterpret_cast<const float4*>(stg + (r4 + sub) * 36 + col);
                        *reinterpret_cast<float4*>(dst + (size_t)(r4 + sub) * g.N + col) = o;
                    }
                    __syncwarp();
                } else {
                    __align__(16) __nv_bfloat16 h[32];
                    if (EPI == EPI_FWD) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float x = __uint_as_float(v[j]);
                            if (g.bias) x += __ldg(g.bias + n0 + j);
                            if (g.relu) x = fmaxf(x, 0.f);
                            h[j] = __float2bfloat16_rn(x);
                        }
                    } else {
                        __align__(16) uint4 gt[4];
                        if (g.gate) {
                            const uint4* gp = reinterpret_cast<const uint4*>(g.gate + (size_t)m * g.N + n0);
#pragma unroll
                            for (int q = 0; q < 4; ++q) gt[q] = __ldg(gp + q);
                        }
                        const __nv_bfloat16* gv = reinterpret_cast<const __nv_bfloat16*>(gt);
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float x = __uint_as_float(v[j]);
                            if (g.gate && !(__bfloat162float(gv[j]) > 0.f)) x = 0.f;
                            h[j] = __float2bfloat16_rn(x);
                        }
                    }
                    uint4* dst = reinterpret_cast<uint4*>(g.out + (size_t)m * g.N + n0);
                    const uint4* src = reinterpret_cast<const uint4*>(h);
#pragma unroll
                    for (int q = 0; q < 4; ++q) dst[q] = src[q];
                    if (g.outT) {  // lanes hold consecutive rows m: a column is 64 contiguous bytes per warp in outT
#pragma unroll
                        for (int j = 0; j < 32; ++j) g.outT[(size_t)(n0 + j) * g.M + m] = h[j];
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
        }
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

}  // namespace gemm
}  // namespace b200ps
