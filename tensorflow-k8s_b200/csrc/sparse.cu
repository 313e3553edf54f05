// sparse.cu -- embedding tables on the PS: gather (pull) and deterministic sorted-segment
// scatter-add + apply (push), include/b200ps.h "Sparse group"; SURVEY.md 8(a) rows a10..a12.
//
// Reference: none -- the reference states only the PS / WORKER roles (k8s-operator.md:6); in a
// TFJob this is TensorFlow's embedding_lookup over a partitioned variable and its sparse
// optimizer ops on the PS tasks.
//
// Rows are partitioned over S shards by TF "div" (the first rows % S shards hold one row more);
// shard s lives on GPU s mod G.  Everything is addressed through per-shard / per-worker base
// pointers that may be peer-mapped (ipc.cu), so routing is not a separate exchange step: the
// kernels of one GPU read all workers' ids and gradient rows wherever they live.
//
// push on one GPU, all device-side (no host round trip; counts stay in the workspace header):
//   1. + 2. stable LSD radix sort of the (worker, position)-ordered ids, 8 bits per pass,
//      ceil(log2(rows)/8) passes, payload = source entry index -> order (id, worker, position);
//      the first pass reads the workers' ids and keeps only those of this GPU's shards.  One
//      pre-pass counts the digits of all passes; each pass is then ONE kernel whose tiles find
//      their offsets by decoupled look-back over the tiles before them
//   3. heads + scan -> segment ids, unique ids, segment offsets (one single-pass scan)
//   4. one lane group per unique id folds its gradient rows in that fixed order (chunks of 256,
//      see oracle/ps_oracle.c segment_sum), scales by 1/W and updates the row (SGD / Adagrad)
// Integer outputs are bit-exact against the oracle; the fold order makes the rows bit-exact too.
//
// HBM-bound integer/byte work: ids 8 B, rows dim*4 B; sort passes are overhead traffic
// (12 B per kept id per pass).  No tensor cores.
#include <cuda_runtime.h>

#include <cstring>
#include <vector>

#include "common.h"
#include "cuda_util.h"

namespace b200ps {
namespace {

// Build-time tuning knobs (-DB200PS_...) with the values measured best on B200 for C2 (2.1 M ids per push):
// keys per thread 8 (4 / 12 / 16: within 10 %), look-back window 8 tiles (16: +9 %), 4 resident CTAs per SM
// for the radix passes, 1024-thread CTAs for the segment scan (256 / 512: +5 / +4 %) -- profiles/r2_summary.md.
constexpr int TILE_THREADS = 256;
#ifndef B200PS_SWEEP_ITEMS
#define B200PS_SWEEP_ITEMS 8
#endif
constexpr int ITEMS = B200PS_SWEEP_ITEMS;
constexpr int TILE = TILE_THREADS * ITEMS;  // entries per CTA in the partition / radix passes
constexpr int WARPS = TILE_THREADS / 32;
constexpr int BINS = 256;
constexpr int MAX_PASSES = 4;  // 32-bit keys, 8 bits per pass
#ifndef B200PS_SWEEP_LOOKBACK
#define B200PS_SWEEP_LOOKBACK 8
#endif
#ifndef B200PS_SWEEP_MINB
#define B200PS_SWEEP_MINB 4
#endif
constexpr int LOOKBACK = B200PS_SWEEP_LOOKBACK;  // predecessor tiles a radix pass fetches per look-back step
#ifndef B200PS_SCAN_THREADS
#define B200PS_SCAN_THREADS 1024
#endif
constexpr int SCAN_THREADS = B200PS_SCAN_THREADS;
constexpr int SCAN_TILE = SCAN_THREADS * 4;
constexpr uint32_t SEG_CHUNK = 256;  // == PS_SEG_CHUNK of the oracle
constexpr uint32_t NONE = 0xFFFFFFFFu;

enum { HDR_KEPT = 0, HDR_UNIQUE = 1, HDR_BAD = 2, HDR_NLONG = 4, HDR_NCHUNKS = 5,
       HDR_TAG = 8,       // identity of the plan (entries, geometry): written by sort_segment, checked by apply
       HDR_MISMATCH = 9,  // raised by an apply that was given another push's workspace
       HDR_STUCK = 10,    // a look-back gave up waiting (never expected: a tile only waits for tiles already running)
       HDR_TICKET = 16,   // tile counters of the radix passes: HDR_TICKET + pass
       HDR_WORDS = 64 };

// Unsigned 32-bit division by a run-time constant without a divide instruction (Granlund &
// Montgomery): exact for every 32-bit numerator.  A hardware-free u64 divide costs ~100
// instructions per lane and made the row kernels issue-bound (ncu: issue slots 48 % busy).
struct FastDiv {
    uint32_t d, magic, shift;
    __host__ void set(uint32_t divisor) {
        d = divisor ? divisor : 1;
        uint32_t L = 0;
        while (L < 32 && (1ull << L) < d) ++L;  // ceil(log2 d)
        if ((1ull << L) == d) {
            magic = 0;  // power of two: a shift
            shift = L;
        } else {
            magic = (uint32_t)((((1ull << L) - d) << 32) / d + 1);
            shift = L - 1;
        }
    }
    __device__ __forceinline__ uint32_t div(uint32_t n) const {
        if (magic == 0) return n >> shift;
        const uint32_t t = __umulhi(n, magic);
        return (t + ((n - t) >> 1)) >> shift;
    }
};

struct Geom {
    uint64_t rows;            // < 2^32
    uint32_t q, r, t;         // div partition: q = rows / S, r = rows % S, t = r * (q + 1)
    FastDiv by_q1, by_q, by_world;  // / (q + 1), / q, / world
    int dim, nshards, rank, world;
};

struct IdSources {
    const int64_t* ids[B200PS_MAX_WORKERS];
    uint32_t per_worker;
    uint32_t n_total;
    FastDiv by_per_worker;
};

struct RowSources {
    const float* rows[B200PS_MAX_WORKERS];
    uint32_t per_worker;
    FastDiv by_per_worker;
};

struct ShardBases {
    float* base[B200PS_MAX_WORKERS];   // row 0 of shard s (table)
    float* accum[B200PS_MAX_WORKERS];  // row 0 of shard s (Adagrad accumulator) or null
};

// owner shard and row inside it (ids are < rows < 2^32, so 32-bit arithmetic is exact)
__device__ __forceinline__ int owner_of(const Geom& g, uint64_t id64, uint64_t* local_row) {
    const uint32_t id = (uint32_t)id64;
    if (id < g.t) {
        const uint32_t s = g.by_q1.div(id);
        *local_row = id - s * (g.q + 1);
        return (int)s;
    }
    const uint32_t k = g.by_q.div(id - g.t);
    *local_row = (id - g.t) - k * g.q;
    return (int)(g.r + k);
}

// does shard s live on this GPU?  (shard s -> GPU s mod world)
__device__ __forceinline__ int gpu_of(const Geom& g, int s) { return s - (int)(g.by_world.div((uint32_t)s) * (uint32_t)g.world); }

// ------------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------------
struct Workspace {
    uint32_t* hdr;
    uint32_t* ghist;      // MAX_PASSES x BINS digit counts of the kept keys (pre-pass), zeroed with the header
    uint64_t* lb;         // look-back state of the segment scan (tile counter + one word per tile), zeroed with the header
    uint64_t* sweep;      // nb x BINS words: (tile, digit) look-back state of the radix passes, zeroed with the header
    size_t zero_bytes;    // hdr .. end of sweep
    uint32_t* key[2];
    uint32_t* val[2];
    uint32_t* seg;        // n + 1: head-count scan, then segment ids
    uint32_t* unique;     // n
    uint32_t* seg_off;    // n + 1
    // tail, used by the apply step only (hot ids): segments longer than SEG_CHUNK
    uint32_t* long_u;     // unique index of each long segment
    uint32_t* long_base;  // its first chunk in `chunks` / `partials`
    uint2* chunks;        // (unique index, chunk number) of every chunk of every long segment
    float* partials;      // one row of dim floats per chunk
    uint32_t max_long, max_chunks;
    uint32_t nb;
    size_t bytes;
};

size_t align_up(size_t x) { return (x + 255) / 256 * 256; }

// what identifies a plan: the push it was made for (never 0)
uint32_t plan_tag(uint64_t n_total, uint64_t per_worker, const b200ps_table_geom_t* g) {
    uint64_t z = n_total * 0x9E3779B97F4A7C15ull ^ per_worker * 0xBF58476D1CE4E5B9ull ^ g->rows * 0x94D049BB133111EBull ^
                 ((uint64_t)g->nshards << 40) ^ ((uint64_t)g->rank << 20) ^ (uint64_t)g->world;
    z ^= z >> 31;
    z *= 0xD6E8FEB86659FD93ull;
    z ^= z >> 29;
    const uint32_t t = (uint32_t)z;
    return t ? t : 1u;
}

// dim == 0: the part sort_segment and plan_read use (its offsets do not depend on dim)
Workspace carve(void* base, uint64_t n, int dim) {
    Workspace w{};
    char* p = (char*)base;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        void* r = p ? (void*)(p + off) : nullptr;
        off += align_up(bytes);
        return r;
    };
    w.nb = (uint32_t)((n + TILE - 1) / TILE);
    if (w.nb == 0) w.nb = 1;
    w.hdr = (uint32_t*)take(HDR_WORDS * 4);
    w.ghist = (uint32_t*)take((size_t)MAX_PASSES * BINS * 4);
    w.lb = (uint64_t*)take(((n + 1 + SCAN_TILE - 1) / SCAN_TILE + 2) * 8);
    w.sweep = (uint64_t*)take((size_t)w.nb * BINS * 8);
    w.zero_bytes = off;
    for (int i = 0; i < 2; ++i) w.key[i] = (uint32_t*)take(n * 4);
    for (int i = 0; i < 2; ++i) w.val[i] = (uint32_t*)take(n * 4);
    w.seg = (uint32_t*)take((n + 1) * 4);
    w.unique = (uint32_t*)take(n * 4);
    w.seg_off = (uint32_t*)take((n + 1) * 4);
    // a long segment has > SEG_CHUNK entries: at most n/257 of them, fewer than n/128 + 1 chunks in all
    w.max_long = (uint32_t)(n / (SEG_CHUNK + 1) + 1);
    w.max_chunks = (uint32_t)(n / (SEG_CHUNK / 2) + 2);
    w.long_u = (uint32_t*)take((size_t)w.max_long * 4);
    w.long_base = (uint32_t*)take((size_t)w.max_long * 4);
    w.chunks = (uint2*)take((size_t)w.max_chunks * 8);
    w.partials = (float*)take((size_t)w.max_chunks * (size_t)dim * 4);
    w.bytes = off;
    return w;
}

// ------------------------------------------------------------------------------------------------
// block-level exclusive scan of uint32 (building block of the single-pass scan below)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t x, uint32_t* total) {
    __shared__ uint32_t warp_sums[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    uint32_t inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += y;
    }
    __syncthreads();  // protects warp_sums across repeated calls
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < nw ? warp_sums[lane] : 0, si = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, si, d);
            if (lane >= d) si += y;
        }
        warp_sums[lane] = si - s;  // exclusive warp offsets
        if (lane == 31) warp_sums[31] = si - s, *total = si;
    }
    __syncthreads();
    return inc - x + warp_sums[wid];
}

// ------------------------------------------------------------------------------------------------
// Segment pass: heads + scan + segments in ONE launch (single-pass scan with decoupled look-back).
// The scanned quantity is the head flag of every sorted key (key differs from the key before it), so
// the scan inside a warp is ballots and popcounts: warp w of a tile owns 128 consecutive entries,
// four rounds of 32, lane-consecutive -- every load and store is coalesced.  Every CTA takes the next
// tile number from a counter (so a tile only ever waits for tiles that are already running),
// publishes (flag, aggregate) as ONE 64-bit word, and warp 0 sums its predecessors' words 32 at a
// time until it meets one that carries an inclusive prefix.
// state[0] = tile counter, state[1 + t] = word of tile t: bits 32..33 = 1 aggregate | 2 inclusive.
// ------------------------------------------------------------------------------------------------
struct SegOut {
    const uint32_t* key;
    uint32_t* hdr;
    uint32_t* seg;
    uint32_t* unique;
    uint32_t* seg_off;
    uint32_t tag;
};

__device__ __forceinline__ uint64_t ld_state(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(uint64_t* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

__global__ void __launch_bounds__(SCAN_THREADS) segment_scan_kernel(uint64_t* state, const __grid_constant__ SegOut so) {
    __shared__ uint32_t warp_excl[32];
    __shared__ uint32_t tile_sh, prefix_sh;
    if (threadIdx.x == 0) tile_sh = atomicAdd(reinterpret_cast<unsigned int*>(state), 1u);
    const uint32_t kept = so.hdr[HDR_KEPT];
    __syncthreads();
    const uint32_t tile = tile_sh, ntiles = gridDim.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t wbase = tile * SCAN_TILE + wid * 128u;
    uint32_t kk[4], ball[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t idx = wbase + k * 32 + lane;
        kk[k] = idx < kept ? so.key[idx] : 0u;
    }
    const uint32_t before = (lane == 0 && wbase > 0 && wbase - 1 < kept) ? so.key[wbase - 1] : 0u;
    uint32_t wtotal = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t idx = wbase + k * 32 + lane;
        uint32_t prev = __shfl_up_sync(0xffffffffu, kk[k], 1);
        const uint32_t last = k ? __shfl_sync(0xffffffffu, kk[k ? k - 1 : 0], 31) : before;
        if (lane == 0) prev = last;
        ball[k] = __ballot_sync(0xffffffffu, idx < kept && (idx == 0 || kk[k] != prev));
        wtotal += __popc(ball[k]);
    }
    if (lane == 0) warp_excl[wid] = wtotal;
    __syncthreads();
    if (wid == 0) {  // warp 0: scan the warp totals, publish the tile, look back
        const uint32_t t = lane < SCAN_THREADS / 32 ? warp_excl[lane] : 0u;
        uint32_t inc = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += y;
        }
        warp_excl[lane] = inc - t;
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        uint32_t prefix = 0;
        if (tile == 0) {
            if (lane == 0) st_state(state + 1, (2ull << 32) | total);
        } else {
            if (lane == 0) st_state(state + 1 + tile, (1ull << 32) | total);
            int base = (int)tile - 1;
            for (;;) {
                const int j = base - lane;
                uint64_t v = 2ull << 32;  // before tile 0: an inclusive prefix of 0
                if (j >= 0) do { v = ld_state(state + 1 + j); } while ((v >> 32) == 0);
                const unsigned inc_mask = __ballot_sync(0xffffffffu, (v >> 32) == 2);
                const int stop = inc_mask ? __ffs(inc_mask) - 1 : 31;  // nearest predecessor that carries an inclusive prefix
                uint32_t val = lane <= stop ? (uint32_t)v : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                prefix += val;
                if (inc_mask) break;
                base -= 32;
            }
            if (lane == 0) st_state(state + 1 + tile, (2ull << 32) | (uint32_t)(prefix + total));
        }
        if (lane == 0) {
            prefix_sh = prefix;
            if (tile == ntiles - 1) {
                const uint32_t grand = prefix + total;
                so.hdr[HDR_UNIQUE] = grand;
                so.seg_off[grand] = kept;
                so.hdr[HDR_TAG] = so.tag;  // the plan is complete and belongs to this push
            }
        }
    }
    __syncthreads();
    uint32_t e = prefix_sh + warp_excl[wid];  // segments before this warp's first entry
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t idx = wbase + k * 32 + lane;
        const uint32_t head = (ball[k] >> lane) & 1u;
        const uint32_t sid = e + __popc(ball[k] & lt) + head - 1u;
        if (idx < kept) {
            so.seg[idx] = sid;
            if (head) {
                so.unique[sid] = kk[k];
                so.seg_off[sid] = idx;
            }
        }
        e += __popc(ball[k]);
    }
}

// ------------------------------------------------------------------------------------------------
// Stable LSD radix sort, one kernel per 8-bit pass ("onesweep": Adinets & Merrill).
//   digit_hist_kernel   one pre-pass over the workers' ids: digit counts of the kept keys for ALL
//                       passes (a key's digits do not depend on where earlier passes put it).
//   sweep_pass_kernel   MODE 0 = the first pass: reads the workers' ids directly and keeps only the
//                       entries of this GPU's shards (a stable partition fused into the pass);
//                       MODE 1 = later passes.  Every CTA draws its tile number from a counter, ranks
//                       its 2048 entries, and thread d publishes the tile's count of digit d as one
//                       64-bit word (flag, count) and sums the words of the tiles before it -- LOOKBACK
//                       of them in flight at once -- until one carries an inclusive prefix.  A tile only
//                       waits for tiles with smaller numbers, which are running or done.
// State word: bits 32.. = 2 * pass + 1 (tile aggregate) | 2 * pass + 2 (inclusive prefix); anything
// else (zero, or an earlier pass's flag) is "not there yet", so one region serves all passes.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE_THREADS) digit_hist_kernel(const __grid_constant__ IdSources src,
                                                                  const __grid_constant__ Geom g, uint32_t* hdr,
                                                                  uint32_t* ghist, int passes) {
    __shared__ uint32_t cnt[MAX_PASSES * BINS];
    for (int i = threadIdx.x; i < MAX_PASSES * BINS; i += TILE_THREADS) cnt[i] = 0;
    __syncthreads();
    uint32_t bad = 0;
    constexpr int PER = 8;  // ids in flight per thread
    for (uint64_t i0 = (uint64_t)blockIdx.x * (TILE_THREADS * PER) + threadIdx.x; i0 < src.n_total; i0 += (uint64_t)gridDim.x * (TILE_THREADS * PER)) {
        int64_t id[PER];
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const uint64_t idx = i0 + (uint64_t)k * TILE_THREADS;
            id[k] = 0;
            if (idx < src.n_total) {
                const uint32_t w = src.by_per_worker.div((uint32_t)idx);
                id[k] = src.ids[w][(uint32_t)idx - w * src.per_worker];
            }
        }
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            if (i0 + (uint64_t)k * TILE_THREADS >= src.n_total) continue;
            if (id[k] < 0 || (uint64_t)id[k] >= g.rows) {
                ++bad;
                continue;
            }
            const uint32_t key = (uint32_t)id[k];
            uint64_t lr;
            if (gpu_of(g, owner_of(g, (uint64_t)id[k], &lr)) != g.rank) continue;
            for (int p = 0; p < passes; ++p) atomicAdd(&cnt[p * BINS + ((key >> (8 * p)) & 255u)], 1u);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(&hdr[HDR_BAD], bad);
    __syncthreads();
    for (int i = threadIdx.x; i < passes * BINS; i += TILE_THREADS)
        if (cnt[i]) atomicAdd(&ghist[i], cnt[i]);
}

struct SweepArgs {
    const uint32_t* key_in;
    const uint32_t* val_in;
    uint32_t* key_out;
    uint32_t* val_out;
    const uint32_t* ghist;  // this pass's BINS digit counts
    uint64_t* state;        // [tile][BINS]
    uint32_t* hdr;
    int pass;
    int shift;
};

template <int MODE>
__global__ void __launch_bounds__(TILE_THREADS, B200PS_SWEEP_MINB) sweep_pass_kernel(const __grid_constant__ SweepArgs a,
                                                                  const __grid_constant__ IdSources src,
                                                                  const __grid_constant__ Geom g) {
    __shared__ uint32_t cnt[WARPS][BINS];
    __shared__ uint32_t tile_start[BINS];  // first staged position of each bin
    __shared__ uint32_t gbase[BINS];       // first global position of (bin, this tile)
    __shared__ uint32_t skey[TILE], sval[TILE];
    __shared__ uint32_t tile_sh, kept_sh, tile_total;
    for (int i = threadIdx.x; i < WARPS * BINS; i += TILE_THREADS) (&cnt[0][0])[i] = 0;
    if (threadIdx.x == 0) tile_sh = atomicAdd(&a.hdr[HDR_TICKET + a.pass], 1u);
    // first output position of each digit = exclusive scan of the digit counts; their sum = #kept
    const uint32_t dbase = block_exclusive_scan(a.ghist[threadIdx.x], &kept_sh);  // ends with a barrier
    const uint32_t tile = tile_sh, kept = kept_sh;
    const uint32_t limit = MODE == 0 ? src.n_total : kept;
    if (MODE == 0 && tile == 0 && threadIdx.x == 0) a.hdr[HDR_KEPT] = kept;
    if ((uint64_t)tile * TILE >= limit) return;  // nothing here, and no tile with entries comes after this one
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    // warp w owns entries [w*256, w*256+256) of the tile, 32 consecutive entries per batch: the
    // (batch, lane) order is the input order, which makes the ranks stable
    const uint32_t base = tile * TILE + w * (32 * ITEMS);
    uint32_t key[ITEMS], val[ITEMS], bin[ITEMS], rank[ITEMS];
    // all of the thread's entries are fetched before the first ranking step: the warp-synchronous
    // ranking below orders memory operations, so loads issued inside it would complete one by one
    if (MODE == 0) {
        int64_t id[ITEMS];
#pragma unroll
        for (int b = 0; b < ITEMS; ++b) {
            const uint32_t idx = base + b * 32 + lane;
            id[b] = -1;
            if (idx < limit) {
                const uint32_t sw = src.by_per_worker.div(idx);
                id[b] = src.ids[sw][idx - sw * src.per_worker];
            }
        }
#pragma unroll
        for (int b = 0; b < ITEMS; ++b) {
            val[b] = base + b * 32 + lane;
            key[b] = (uint32_t)id[b];
            bin[b] = NONE;
            uint64_t lr;
            if (id[b] >= 0 && (uint64_t)id[b] < g.rows && gpu_of(g, owner_of(g, (uint64_t)id[b], &lr)) == g.rank) bin[b] = key[b] & 255u;
        }
    } else {
#pragma unroll
        for (int b = 0; b < ITEMS; ++b) {
            const uint32_t idx = base + b * 32 + lane;
            key[b] = val[b] = 0;
            if (idx < limit) {
                key[b] = a.key_in[idx];
                val[b] = a.val_in[idx];
            }
            bin[b] = idx < limit ? (key[b] >> a.shift) & 255u : NONE;
        }
    }
#pragma unroll
    for (int b = 0; b < ITEMS; ++b) {
        const uint32_t peers = __match_any_sync(0xffffffffu, bin[b]);
        uint32_t start = 0;
        if (bin[b] != NONE) start = cnt[w][bin[b]];
        __syncwarp();
        if (bin[b] != NONE && (peers & lt) == 0) cnt[w][bin[b]] = start + __popc(peers);
        __syncwarp();
        rank[b] = start + __popc(peers & lt);
    }
    __syncthreads();
    {
        const int t = threadIdx.x;  // one bin per thread: exclusive offsets over the warps
        uint32_t run = 0;
#pragma unroll
        for (int k = 0; k < WARPS; ++k) {
            const uint32_t c = cnt[k][t];
            cnt[k][t] = run;
            run += c;
        }
        // publish this tile's count of digit t, then sum the tiles before it
        uint64_t* const col = a.state + t;
        const uint32_t fa = 2u * (uint32_t)a.pass + 1u, fi = fa + 1u;
        uint32_t excl = 0;
        if (tile == 0) {
            st_state(col, ((uint64_t)fi << 32) | run);
        } else {
            st_state(col + (size_t)tile * BINS, ((uint64_t)fa << 32) | run);
            const uint64_t* win = col + (size_t)(tile - 1) * BINS;  // nearest predecessor of the window
            int64_t left = (int64_t)tile;                           // predecessors from `win` down to tile 0
            uint32_t spins = 0;
            for (;;) {
                uint64_t v[LOOKBACK];
#pragma unroll
                for (int k = 0; k < LOOKBACK; ++k) v[k] = k < left ? ld_state(win - (size_t)k * BINS) : ((uint64_t)fi << 32);
                bool ready = true;
#pragma unroll
                for (int k = 0; k < LOOKBACK; ++k) ready &= ((uint32_t)(v[k] >> 32) - fa) < 2u;
                if (!ready) {  // a tile of the window has not published yet: fetch the window again
                    if (++spins > (1u << 22)) {  // seconds: give up loudly instead of hanging the GPU
                        a.hdr[HDR_STUCK] = 1u;
                        break;
                    }
                    continue;
                }
                bool done = false;
#pragma unroll
                for (int k = 0; k < LOOKBACK; ++k) {
                    excl += done ? 0u : (uint32_t)v[k];
                    done |= (uint32_t)(v[k] >> 32) == fi;
                }
                if (done) break;
                win -= (size_t)LOOKBACK * BINS;
                left -= LOOKBACK;
            }
            st_state(col + (size_t)tile * BINS, ((uint64_t)fi << 32) | (uint32_t)(excl + run));
        }
        gbase[t] = dbase + excl;
        tile_start[t] = block_exclusive_scan(run, &tile_total);  // over the 256 bins; total -> shared
    }
    __syncthreads();
    // Stage the tile in shared memory in (bin, input order) -- its position inside the tile -- and
    // write it out from there: entries of one bin leave as one contiguous run instead of 32
    // scattered 4-byte stores per warp instruction (ncu: 30 sectors per store request before).
#pragma unroll
    for (int b = 0; b < ITEMS; ++b) {
        if (bin[b] == NONE) continue;
        const uint32_t lp = tile_start[bin[b]] + cnt[w][bin[b]] + rank[b];
        skey[lp] = key[b];
        sval[lp] = val[b];
    }
    __syncthreads();
    for (uint32_t lp = threadIdx.x; lp < tile_total; lp += TILE_THREADS) {
        const uint32_t k = skey[lp];
        const uint32_t bn = (k >> a.shift) & 255u;
        const uint32_t dst = gbase[bn] + (lp - tile_start[bn]);
        a.key_out[dst] = k;
        a.val_out[dst] = sval[lp];
    }
}

// ------------------------------------------------------------------------------------------------
// a12: one lane group per unique id
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg_row4(const float* row, int c) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(reinterpret_cast<const float4*>(row) + c));
    return r;
}

__device__ __forceinline__ float4 add4(const float4& a, const float4& b) {
    return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
}

__device__ __forceinline__ const float* grad_row(const RowSources& gs, uint32_t entry, int dim) {
    const uint32_t w = gs.by_per_worker.div(entry), pos = entry - w * gs.per_worker;
    return gs.rows[w] + (uint64_t)pos * (uint64_t)dim;
}

template <bool ADAGRAD>
__device__ __forceinline__ float upd1(float g_sum, float row, float& acc, float inv_w, float lr) {
    const float g = __fmul_rn(g_sum, inv_w);
    if (!ADAGRAD) return __fsub_rn(row, __fmul_rn(lr, g));
    acc = __fadd_rn(acc, __fmul_rn(g, g));
    return __fsub_rn(row, __fdiv_rn(__fmul_rn(lr, g), __fsqrt_rn(acc)));
}

struct ApplyArgs {
    uint32_t tag;  // plan_tag of the push this apply belongs to
    const uint32_t* hdr_ro;
    uint32_t* hdr;
    const uint32_t* perm;
    const uint32_t* unique;
    const uint32_t* seg_off;
    uint32_t* long_u;
    uint32_t* long_base;
    uint2* chunks;
    float* partials;
    float inv_w, lr;
};

// continues the left fold of column c: acc + row[j0] + row[j0+1] + ... (sorted entries [j0, j1) of
// the plan), F gradient rows in flight, added strictly in order
template <int F = 8>
__device__ __forceinline__ float4 fold_rows(float4 acc, const RowSources& gs, const uint32_t* __restrict__ perm, uint32_t j0,
                                            uint32_t j1, int dim, int c) {
    uint32_t j = j0;
    for (; j + F <= j1; j += F) {
        float4 r[F];
#pragma unroll
        for (int k = 0; k < F; ++k) r[k] = ldg_row4(grad_row(gs, perm[j + k], dim), c);
#pragma unroll
        for (int k = 0; k < F; ++k) acc = add4(acc, r[k]);
    }
    for (; j < j1; ++j) acc = add4(acc, ldg_row4(grad_row(gs, perm[j], dim), c));
    return acc;
}

template <bool ADAGRAD>
__device__ __forceinline__ void update_row(float* row, float* arow, int c, const float4& total, float inv_w, float lr) {
    float4 r = reinterpret_cast<float4*>(row)[c];
    float4 ac = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ADAGRAD) ac = reinterpret_cast<float4*>(arow)[c];
    r.x = upd1<ADAGRAD>(total.x, r.x, ac.x, inv_w, lr);
    r.y = upd1<ADAGRAD>(total.y, r.y, ac.y, inv_w, lr);
    r.z = upd1<ADAGRAD>(total.z, r.z, ac.z, inv_w, lr);
    r.w = upd1<ADAGRAD>(total.w, r.w, ac.w, inv_w, lr);
    reinterpret_cast<float4*>(row)[c] = r;
    if (ADAGRAD) reinterpret_cast<float4*>(arow)[c] = ac;
}

// Segments of at most SEG_CHUNK entries: fold and update here.  Longer ones (hot ids) are only
// registered -- their chunks are folded by all lane groups of long_partials_kernel and finished by
// long_final_kernel, in the chunk order the oracle defines.
//
// Two phases per tile of 256 unique ids (like sparse_serve_kernel).  A: one THREAD per unique id
// reads its segment bounds and id, resolves the table row once and leaves a descriptor in shared
// memory.  B1: segments of one or two entries (99 % of them for uniform ids) -- one LANE GROUP takes
// R unique ids at a time, and the R table rows and their gradient rows are all in flight before the
// first add: this is where the memory parallelism comes from.  B2: the longer segments, one per lane
// group at a time, folded in order with eight gradient rows in flight.
//
// Which ids a tile holds: runs of 8 consecutive unique ids, run j of the table going to tile
// j mod T (T tiles).  Eight neighbours share the 32-byte sectors of seg_off / unique / perm, so
// phase A reads ~20 bytes per id instead of four separate sectors; and sorted neighbours -- under
// Zipf skew the hot ids with their long segments sit together at the front -- are still dealt out
// over all CTAs instead of landing in one.
__device__ __forceinline__ uint32_t tile_uid(uint32_t tile, uint32_t r, uint32_t ntiles) {
    return (tile + (r >> 3) * ntiles) * 8u + (r & 7u);  // < 256 * ntiles < n_unique + 256: the entry count is capped below 2^32 - 2^16
}

#ifndef B200PS_APPLY_ROWS
#define B200PS_APPLY_ROWS 4  // unique ids in flight per lane group (phase B1)
#endif
#ifndef B200PS_APPLY_MINB
#define B200PS_APPLY_MINB 3  // resident CTAs per SM the register allocation must allow (R = 4: 72 / 80 registers, no spills)
#endif
template <int LG, bool ADAGRAD>
__global__ void __launch_bounds__(256, B200PS_APPLY_MINB) sparse_apply_kernel(const __grid_constant__ ShardBases tb,
                                                                              const __grid_constant__ RowSources gs,
                                                                              const __grid_constant__ Geom g,
                                                                              const __grid_constant__ ApplyArgs a) {
    constexpr int GROUPS = 256 / LG;
    constexpr int R = B200PS_APPLY_ROWS;
    __shared__ float* s_row[256];
    __shared__ float* s_arow[ADAGRAD ? 256 : 1];
    __shared__ uint32_t s_s0[256], s_len[256], s_base[256], s_first[256], s_second[256];
    if (a.hdr_ro[HDR_TAG] != a.tag) {  // the workspace holds another push's plan (or none): touch nothing
        if (blockIdx.x == 0 && threadIdx.x == 0) a.hdr[HDR_MISMATCH] = 1u;
        return;
    }
    const uint32_t n_unique = a.hdr_ro[HDR_UNIQUE];
    const int lane = threadIdx.x % LG, grp = threadIdx.x / LG;
    const int dim4 = g.dim / 4;
    const uint32_t ntiles = (n_unique + 255u) / 256u;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        {  // phase A
            const uint32_t u = tile_uid(tile, threadIdx.x, ntiles);
            uint32_t s0 = 0, len = 0, cbase = 0, first = 0, second = 0;  // len 0: past the last unique id
            float* row = nullptr;
            if (u < n_unique) {
                s0 = a.seg_off[u];
                len = a.seg_off[u + 1] - s0;
                first = a.perm[s0];  // source entries of the segment's first rows: one dependent load less in phase B
                if (len >= 2) second = a.perm[s0 + 1];
                uint64_t lrow;
                const int s = owner_of(g, a.unique[u], &lrow);
                row = tb.base[s] + lrow * (uint64_t)g.dim;
                if (ADAGRAD) s_arow[threadIdx.x] = tb.accum[s] + lrow * (uint64_t)g.dim;
                if (len > SEG_CHUNK) {
                    const uint32_t nchunks = (len + SEG_CHUNK - 1) / SEG_CHUNK;
                    const uint32_t idx = atomicAdd(&a.hdr[HDR_NLONG], 1u);
                    cbase = atomicAdd(&a.hdr[HDR_NCHUNKS], nchunks);
                    a.long_u[idx] = u;
                    a.long_base[idx] = cbase;
                }
            }
            s_row[threadIdx.x] = row;
            s_s0[threadIdx.x] = s0;
            s_len[threadIdx.x] = len;
            s_base[threadIdx.x] = cbase;
            s_first[threadIdx.x] = first;
            s_second[threadIdx.x] = second;
        }
        __syncthreads();
        for (int r0 = grp; r0 < 256; r0 += GROUPS * R) {  // phase B1: one or two entries
            for (int c = lane; c < dim4; c += LG) {
                float4 rv[R], ac[R], g0[R], g1[R];
#pragma unroll
                for (int k = 0; k < R; ++k) {
                    const int r = r0 + k * GROUPS;
                    const uint32_t len = r < 256 ? s_len[r] : 0u;
                    rv[k] = ac[k] = g0[k] = g1[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (len - 1u >= 2u) continue;
                    rv[k] = reinterpret_cast<const float4*>(s_row[r])[c];
                    if (ADAGRAD) ac[k] = reinterpret_cast<const float4*>(s_arow[r])[c];
                    g0[k] = ldg_row4(grad_row(gs, s_first[r], g.dim), c);
                    if (len == 2u) g1[k] = ldg_row4(grad_row(gs, s_second[r], g.dim), c);
                }
#pragma unroll
                for (int k = 0; k < R; ++k) {
                    const int r = r0 + k * GROUPS;
                    const uint32_t len = r < 256 ? s_len[r] : 0u;
                    if (len - 1u >= 2u) continue;
                    const float4 acc = len == 2u ? add4(g0[k], g1[k]) : g0[k];
                    rv[k].x = upd1<ADAGRAD>(acc.x, rv[k].x, ac[k].x, a.inv_w, a.lr);
                    rv[k].y = upd1<ADAGRAD>(acc.y, rv[k].y, ac[k].y, a.inv_w, a.lr);
                    rv[k].z = upd1<ADAGRAD>(acc.z, rv[k].z, ac[k].z, a.inv_w, a.lr);
                    rv[k].w = upd1<ADAGRAD>(acc.w, rv[k].w, ac[k].w, a.inv_w, a.lr);
                    reinterpret_cast<float4*>(s_row[r])[c] = rv[k];
                    if (ADAGRAD) reinterpret_cast<float4*>(s_arow[r])[c] = ac[k];
                }
            }
        }
        for (int r = grp; r < 256; r += GROUPS) {  // phase B2: three entries and more
            const uint32_t s0 = s_s0[r], len = s_len[r];
            if (len < 3u) continue;
            if (len > SEG_CHUNK) {  // a hot id: register its chunks, the long kernels do the rest
                const uint32_t nchunks = (len + SEG_CHUNK - 1) / SEG_CHUNK, cbase = s_base[r];
                for (uint32_t jj = lane; jj < nchunks; jj += LG) a.chunks[cbase + jj] = make_uint2(tile_uid(tile, (uint32_t)r, ntiles), jj);
                continue;
            }
            float* row = s_row[r];
            float* arow = ADAGRAD ? s_arow[r] : nullptr;
            for (int c = lane; c < dim4; c += LG) {
                // the row being updated does not depend on the fold: get it in flight first
                float4 rv = reinterpret_cast<const float4*>(row)[c];
                float4 ac = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ADAGRAD) ac = reinterpret_cast<const float4*>(arow)[c];
                float4 acc = ldg_row4(grad_row(gs, s_first[r], g.dim), c);
                acc = fold_rows(acc, gs, a.perm, s0 + 1, s0 + len, g.dim, c);
                rv.x = upd1<ADAGRAD>(acc.x, rv.x, ac.x, a.inv_w, a.lr);
                rv.y = upd1<ADAGRAD>(acc.y, rv.y, ac.y, a.inv_w, a.lr);
                rv.z = upd1<ADAGRAD>(acc.z, rv.z, ac.z, a.inv_w, a.lr);
                rv.w = upd1<ADAGRAD>(acc.w, rv.w, ac.w, a.inv_w, a.lr);
                reinterpret_cast<float4*>(row)[c] = rv;
                if (ADAGRAD) reinterpret_cast<float4*>(arow)[c] = ac;
            }
        }
        __syncthreads();
    }
}

template <int LG>
__global__ void __launch_bounds__(256) long_partials_kernel(const __grid_constant__ RowSources gs, const __grid_constant__ Geom g,
                                                            const __grid_constant__ ApplyArgs a) {
    const uint32_t nchunks = a.hdr_ro[HDR_NCHUNKS];
    const int lane = threadIdx.x % LG;
    const uint32_t group = (blockIdx.x * blockDim.x + threadIdx.x) / LG, ngroups = gridDim.x * blockDim.x / LG;
    const int dim4 = g.dim / 4;
    for (uint32_t ci = group; ci < nchunks; ci += ngroups) {
        const uint2 d = a.chunks[ci];
        const uint32_t s0 = a.seg_off[d.x], s1 = a.seg_off[d.x + 1];
        const uint32_t j0 = s0 + d.y * SEG_CHUNK, j1 = s1 - j0 < SEG_CHUNK ? s1 : j0 + SEG_CHUNK;
        for (int c = lane; c < dim4; c += LG)
            reinterpret_cast<float4*>(a.partials + (uint64_t)ci * (uint64_t)g.dim)[c] =
                fold_rows<16>(ldg_row4(grad_row(gs, a.perm[j0], g.dim), c), gs, a.perm, j0 + 1, j1, g.dim, c);
    }
}

template <int LG, bool ADAGRAD>
__global__ void __launch_bounds__(256) long_final_kernel(const __grid_constant__ ShardBases tb, const __grid_constant__ Geom g,
                                                         const __grid_constant__ ApplyArgs a) {
    const uint32_t nlong = a.hdr_ro[HDR_NLONG];
    const int lane = threadIdx.x % LG;
    const uint32_t group = (blockIdx.x * blockDim.x + threadIdx.x) / LG, ngroups = gridDim.x * blockDim.x / LG;
    const int dim4 = g.dim / 4;
    for (uint32_t li = group; li < nlong; li += ngroups) {
        const uint32_t u = a.long_u[li], base = a.long_base[li];
        const uint32_t len = a.seg_off[u + 1] - a.seg_off[u], nchunks = (len + SEG_CHUNK - 1) / SEG_CHUNK;
        uint64_t lrow;
        const int s = owner_of(g, a.unique[u], &lrow);
        float* row = tb.base[s] + lrow * (uint64_t)g.dim;
        float* arow = ADAGRAD ? tb.accum[s] + lrow * (uint64_t)g.dim : nullptr;
        for (int c = lane; c < dim4; c += LG) {
            const float4* p = reinterpret_cast<const float4*>(a.partials + (uint64_t)base * (uint64_t)g.dim) + c;
            float4 total = p[0];
            uint32_t j = 1;
            for (; j + 8 <= nchunks; j += 8) {  // eight partial rows in flight, added in chunk order
                float4 r[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) r[k] = p[(uint64_t)(j + k) * (uint64_t)dim4];
#pragma unroll
                for (int k = 0; k < 8; ++k) total = add4(total, r[k]);
            }
            for (; j < nchunks; ++j) total = add4(total, p[(uint64_t)j * (uint64_t)dim4]);
            update_row<ADAGRAD>(row, arow, c, total, a.inv_w, a.lr);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// a10: gather (single-worker pull; the all-worker owner-side form is sparse_serve_kernel below)
// ------------------------------------------------------------------------------------------------
// Copies the rows described in shared memory: entry k of the tile goes from s_src[k] to s_dst[k]
// (null dst: not this GPU's business; null src with a dst: an out-of-range id, answered with zeros).
// One lane group per row, R rows in flight per group.
template <int LG, int R>
__device__ __forceinline__ void copy_tile_rows(const float4* const* s_src, float4* const* s_dst, int count, int dim4) {
    constexpr int GROUPS = 256 / LG;
    const int lane = threadIdx.x % LG, grp = threadIdx.x / LG;
    for (int r0 = grp; r0 < count; r0 += GROUPS * R) {
        const float4* rd[R];
        float4* wr[R];
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const int r = r0 + k * GROUPS;
            rd[k] = r < count ? s_src[r] : nullptr;
            wr[k] = r < count ? s_dst[r] : nullptr;
        }
        for (int c = lane; c < dim4; c += LG) {
            float4 v[R];
#pragma unroll
            for (int k = 0; k < R; ++k) v[k] = rd[k] ? rd[k][c] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < R; ++k)
                if (wr[k]) wr[k][c] = v[k];
        }
    }
}

// Tiles of 256 ids, two phases (as in sparse_serve_kernel): one thread per id resolves the source
// row once into shared memory, then one lane group per row copies it.
template <int LG>
__global__ void __launch_bounds__(256) sparse_gather_kernel(float* __restrict__ out, const __grid_constant__ ShardBases tb,
                                                            const __grid_constant__ Geom g, const int64_t* __restrict__ ids,
                                                            uint64_t n) {
    __shared__ const float4* s_src[256];
    __shared__ float4* s_dst[256];
    const int dim4 = g.dim / 4;
    for (uint64_t base = (uint64_t)blockIdx.x * 256u; base < n; base += (uint64_t)gridDim.x * 256u) {
        const uint64_t k = base + threadIdx.x;
        const float4* rd = nullptr;
        float4* wr = nullptr;
        if (k < n) {
            const int64_t id = ids[k];
            wr = reinterpret_cast<float4*>(out + k * (uint64_t)g.dim);
            if (id >= 0 && (uint64_t)id < g.rows) {  // an out-of-range id keeps rd null: a row of zeros
                uint64_t lrow;
                const int s = owner_of(g, (uint64_t)id, &lrow);
                rd = reinterpret_cast<const float4*>(tb.base[s] + lrow * (uint64_t)g.dim);
            }
        }
        s_src[threadIdx.x] = rd;
        s_dst[threadIdx.x] = wr;
        __syncthreads();
        const uint64_t left = n - base;
        copy_tile_rows<LG, 4>(s_src, s_dst, left < 256u ? (int)left : 256, dim4);
        __syncthreads();
    }
}

// grid for a grid-stride kernel: exactly the CTAs that are resident at once (SMs x occupancy), so
// there is no partial last wave; capped by the work available
template <typename Kern>
int resident_grid(Kern kern, int threads, uint64_t want_blocks) {
    int sms = 0, per_sm = 0;
    if (sm_count(&sms) != cudaSuccess) sms = 148;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 2;
    const uint64_t cap = (uint64_t)sms * (uint64_t)per_sm;
    return (int)(want_blocks < cap ? (want_blocks ? want_blocks : 1) : cap);
}

int lane_group(int dim) {
    const int d4 = dim / 4;
    int lg = 4;
    while (lg < d4 && lg < 32) lg <<= 1;
    return lg;
}

int make_geom(const b200ps_table_geom_t* in, Geom* g) {
    B200PS_REQUIRE(in, B200PS_ERR_INVALID, "null table geometry");
    B200PS_REQUIRE(in->rows >= 1 && in->rows <= 0xFFFFFFFFull, B200PS_ERR_RANGE, "rows must be in 1..2^32-1");
    B200PS_REQUIRE(in->dim >= 4 && in->dim % 4 == 0 && in->dim <= 4096, B200PS_ERR_RANGE,
                   "dim must be a multiple of 4 in 4..4096");
    B200PS_REQUIRE(in->nshards >= 1 && in->nshards <= B200PS_MAX_WORKERS, B200PS_ERR_RANGE, "nshards must be in 1..%d",
                   B200PS_MAX_WORKERS);
    B200PS_REQUIRE(in->world >= 1 && in->rank >= 0 && in->rank < in->world, B200PS_ERR_RANGE, "rank %d of world %d", in->rank,
                   in->world);
    g->rows = in->rows;
    g->q = (uint32_t)(in->rows / (uint64_t)in->nshards);
    g->r = (uint32_t)(in->rows % (uint64_t)in->nshards);
    g->t = g->r * (g->q + 1);
    g->by_q1.set(g->q + 1);
    g->by_q.set(g->q);  // q == 0 only when every id is below t, so the divisor is never used then
    g->by_world.set((uint32_t)in->world);
    g->dim = in->dim;
    g->nshards = in->nshards;
    g->rank = in->rank;
    g->world = in->world;
    return B200PS_OK;
}

int key_passes(uint64_t rows) {
    int bits = 1;
    while (bits < 32 && (1ull << bits) < rows) ++bits;
    return (bits + 7) / 8;
}

int fill_bases(const b200ps_table_geom_t* geom, float* const* table, float* const* accum, ShardBases* tb) {
    B200PS_REQUIRE(table, B200PS_ERR_INVALID, "null shard base list");
    memset(tb, 0, sizeof *tb);
    for (int s = 0; s < geom->nshards; ++s) {
        tb->base[s] = table[s];
        tb->accum[s] = accum ? accum[s] : nullptr;
        if (s % geom->world == geom->rank) {
            B200PS_REQUIRE(table[s] && ((uintptr_t)table[s] & 15) == 0, B200PS_ERR_INVALID,
                           "table shard %d must be a 16-byte aligned device pointer", s);
            if (accum) B200PS_REQUIRE(accum[s] && ((uintptr_t)accum[s] & 15) == 0, B200PS_ERR_INVALID,
                                      "accumulator shard %d must be a 16-byte aligned device pointer", s);
        }
    }
    return B200PS_OK;
}

template <bool ADAGRAD>
int sparse_apply(float* const* d_shard_base, float* const* d_accum_base, const float* const* d_grad_rows, int W,
                 uint64_t ids_per_worker, const b200ps_table_geom_t* geom, float lr, const void* d_workspace, void* stream) {
    Geom g;
    if (int rc = make_geom(geom, &g)) return rc;
    B200PS_REQUIRE(W >= 1 && W <= B200PS_MAX_WORKERS && d_grad_rows && d_workspace, B200PS_ERR_INVALID,
                   "bad worker count or null argument");
    const uint64_t n = (uint64_t)W * ids_per_worker;
    B200PS_REQUIRE(n < 0xFFFF0000ull, B200PS_ERR_RANGE, "W * ids_per_worker must be below 2^32 - 2^16");
    if (n == 0) return B200PS_OK;
    ShardBases tb;
    if (int rc = fill_bases(geom, d_shard_base, d_accum_base, &tb)) return rc;
    RowSources gs{};
    gs.per_worker = (uint32_t)ids_per_worker;
    gs.by_per_worker.set(gs.per_worker);
    for (int w = 0; w < W; ++w) {
        B200PS_REQUIRE(d_grad_rows[w] && ((uintptr_t)d_grad_rows[w] & 15) == 0, B200PS_ERR_INVALID,
                       "gradient rows of worker %d must be a 16-byte aligned device pointer", w);
        gs.rows[w] = d_grad_rows[w];
    }
    const Workspace ws = carve(const_cast<void*>(d_workspace), n, g.dim);
    const int passes = key_passes(g.rows);
    ApplyArgs a{};
    a.tag = plan_tag(n, ids_per_worker, geom);
    a.hdr_ro = ws.hdr;
    a.hdr = ws.hdr;
    a.perm = ws.val[(passes - 1) & 1];  // the first pass wrote buffer 0, each later pass flips
    a.unique = ws.unique;
    a.seg_off = ws.seg_off;
    a.long_u = ws.long_u;
    a.long_base = ws.long_base;
    a.chunks = ws.chunks;
    a.partials = ws.partials;
    a.inv_w = 1.0f / (float)W;
    a.lr = lr;
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const int lg = lane_group(g.dim);
    const uint64_t want = (n * (uint64_t)lg + 255) / 256;
    cudaStream_t st = (cudaStream_t)stream;
    // the long-segment counters belong to this apply (a plan may be applied more than once)
    B200PS_CUDA(cudaMemsetAsync(ws.hdr + HDR_NLONG, 0, 2 * sizeof(uint32_t), st));
#define LAUNCH_APPLY(LG)                                                      \
    sparse_apply_kernel<LG, ADAGRAD><<<resident_grid(sparse_apply_kernel<LG, ADAGRAD>, 256, (n + 255) / 256), 256, 0, st>>>(tb, gs, g, a); \
    long_partials_kernel<LG><<<resident_grid(long_partials_kernel<LG>, 256, want), 256, 0, st>>>(gs, g, a);              \
    long_final_kernel<LG, ADAGRAD><<<sms, 256, 0, st>>>(tb, g, a)
    switch (lg) {
        case 4: LAUNCH_APPLY(4); break;
        case 8: LAUNCH_APPLY(8); break;
        case 16: LAUNCH_APPLY(16); break;
        default: LAUNCH_APPLY(32); break;
    }
#undef LAUNCH_APPLY
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(3, std::memory_order_relaxed);
    return B200PS_OK;
}

}  // namespace
}  // namespace b200ps

using namespace b200ps;

extern "C" size_t b200ps_sparse_workspace_bytes(uint64_t n_total, int dim) {
    return carve(nullptr, n_total ? n_total : 1, dim > 0 ? dim : 0).bytes;
}

extern "C" int b200ps_sparse_gather(float* d_out, const float* const* d_shard_base, const b200ps_table_geom_t* geom,
                                    const int64_t* d_ids, uint64_t n, void* stream) {
    Geom g;
    if (int rc = make_geom(geom, &g)) return rc;
    B200PS_REQUIRE((d_out && d_ids) || n == 0, B200PS_ERR_INVALID, "null output or id pointer");
    B200PS_REQUIRE(d_shard_base, B200PS_ERR_INVALID, "null shard base list");
    if (n == 0) return B200PS_OK;
    B200PS_REQUIRE(((uintptr_t)d_out & 15) == 0, B200PS_ERR_INVALID, "output must be 16-byte aligned");
    ShardBases tb{};
    for (int s = 0; s < g.nshards; ++s) {
        B200PS_REQUIRE(d_shard_base[s] && ((uintptr_t)d_shard_base[s] & 15) == 0, B200PS_ERR_INVALID,
                       "table shard %d must be a 16-byte aligned device pointer", s);
        tb.base[s] = const_cast<float*>(d_shard_base[s]);
    }
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const int lg = lane_group(g.dim);
    cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH_GATHER(LG) \
    sparse_gather_kernel<LG><<<resident_grid(sparse_gather_kernel<LG>, 256, (n + 255) / 256), 256, 0, st>>>(d_out, tb, g, d_ids, n)
    switch (lg) {
        case 4: LAUNCH_GATHER(4); break;
        case 8: LAUNCH_GATHER(8); break;
        case 16: LAUNCH_GATHER(16); break;
        default: LAUNCH_GATHER(32); break;
    }
#undef LAUNCH_GATHER
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_sparse_sort_segment(const int64_t* const* d_ids, int W, uint64_t ids_per_worker,
                                          const b200ps_table_geom_t* geom, void* d_workspace, size_t workspace_bytes,
                                          void* stream) {
    Geom g;
    if (int rc = make_geom(geom, &g)) return rc;
    B200PS_REQUIRE(W >= 1 && W <= B200PS_MAX_WORKERS && d_ids && d_workspace, B200PS_ERR_INVALID,
                   "bad worker count or null argument");
    const uint64_t n = (uint64_t)W * ids_per_worker;
    B200PS_REQUIRE(n >= 1 && n < 0xFFFF0000ull, B200PS_ERR_RANGE, "W * ids_per_worker must be in 1..2^32-2^16-1");
    B200PS_REQUIRE(((uintptr_t)d_workspace & 255) == 0, B200PS_ERR_INVALID, "workspace must be 256-byte aligned");
    const Workspace ws = carve(d_workspace, n, g.dim);
    B200PS_REQUIRE(workspace_bytes >= ws.bytes, B200PS_ERR_RANGE, "workspace of %zu bytes is too small: %zu needed",
                   workspace_bytes, ws.bytes);
    IdSources src{};
    src.per_worker = (uint32_t)ids_per_worker;
    src.by_per_worker.set(src.per_worker);
    src.n_total = (uint32_t)n;
    for (int w = 0; w < W; ++w) {
        B200PS_REQUIRE(d_ids[w], B200PS_ERR_INVALID, "ids of worker %d are null", w);
        src.ids[w] = d_ids[w];
    }
    cudaStream_t st = (cudaStream_t)stream;
    int launches = 0;
    // header counters, digit counts and all look-back state (contiguous), zeroed together
    B200PS_CUDA(cudaMemsetAsync(ws.hdr, 0, ws.zero_bytes, st));

    // 0. digit counts of the kept keys, all passes at once
    const int passes = key_passes(g.rows);
    {
        int sms = 0;
        B200PS_CUDA(sm_count(&sms));
        const uint32_t want = (uint32_t)((n + TILE_THREADS * 8 - 1) / (TILE_THREADS * 8));  // 8 entries per thread and iteration
        const uint32_t grid = want < (uint32_t)(4 * sms) ? want : (uint32_t)(4 * sms);
        digit_hist_kernel<<<grid, TILE_THREADS, 0, st>>>(src, g, ws.hdr, ws.ghist, passes);
        launches += 1;
    }
    // 1 + 2. LSD radix passes, ping-pong between the two buffers; the first one reads the workers'
    // ids and keeps this GPU's entries (key = id, value = source entry w*B + pos)
    SweepArgs a{};
    a.state = ws.sweep;
    a.hdr = ws.hdr;
    int cur = 0;
    for (int p = 0; p < passes; ++p) {
        a.key_in = ws.key[cur];
        a.val_in = ws.val[cur];
        a.key_out = ws.key[p == 0 ? 0 : cur ^ 1];
        a.val_out = ws.val[p == 0 ? 0 : cur ^ 1];
        a.ghist = ws.ghist + p * BINS;
        a.pass = p;
        a.shift = 8 * p;
        if (p == 0) sweep_pass_kernel<0><<<ws.nb, TILE_THREADS, 0, st>>>(a, src, g);
        else sweep_pass_kernel<1><<<ws.nb, TILE_THREADS, 0, st>>>(a, src, g);
        launches += 1;
        if (p > 0) cur ^= 1;
    }
    // 3. segment heads, their scan, segment ids / unique ids / offsets: one launch
    {
        SegOut so{ws.key[cur], ws.hdr, ws.seg, ws.unique, ws.seg_off, plan_tag(n, ids_per_worker, geom)};
        segment_scan_kernel<<<((uint32_t)n + SCAN_TILE - 1) / SCAN_TILE, SCAN_THREADS, 0, st>>>(ws.lb, so);
        launches += 1;
    }
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add((uint64_t)launches, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_sparse_apply_sgd(float* const* d_shard_base, const float* const* d_grad_rows, int W,
                                       uint64_t ids_per_worker, const b200ps_table_geom_t* geom, float lr,
                                       const void* d_workspace, void* stream) {
    return sparse_apply<false>(d_shard_base, nullptr, d_grad_rows, W, ids_per_worker, geom, lr, d_workspace, stream);
}

extern "C" int b200ps_sparse_apply_adagrad(float* const* d_shard_base, float* const* d_accum_base,
                                           const float* const* d_grad_rows, int W, uint64_t ids_per_worker,
                                           const b200ps_table_geom_t* geom, float lr, const void* d_workspace, void* stream) {
    B200PS_REQUIRE(d_accum_base, B200PS_ERR_INVALID, "Adagrad needs the accumulator shards");
    return sparse_apply<true>(d_shard_base, d_accum_base, d_grad_rows, W, ids_per_worker, geom, lr, d_workspace, stream);
}

extern "C" int b200ps_sparse_plan_read(const void* d_workspace, uint64_t n_total, uint64_t rows, uint64_t* n_kept,
                                       uint64_t* n_unique, uint64_t* n_bad, int64_t* h_sorted_ids, uint32_t* h_perm,
                                       int32_t* h_seg_id, int64_t* h_unique_ids, uint32_t* h_seg_offset, void* stream) {
    B200PS_REQUIRE(d_workspace && n_total >= 1 && rows >= 1, B200PS_ERR_INVALID, "null workspace or empty plan");
    cudaStream_t st = (cudaStream_t)stream;
    const Workspace ws = carve(const_cast<void*>(d_workspace), n_total, 0);
    uint32_t hdr[HDR_WORDS];
    B200PS_CUDA(cudaMemcpyAsync(hdr, ws.hdr, sizeof hdr, cudaMemcpyDeviceToHost, st));
    B200PS_CUDA(cudaStreamSynchronize(st));
    B200PS_REQUIRE(hdr[HDR_MISMATCH] == 0, B200PS_ERR_STATE,
                   "an apply call was given this workspace with another push's (W, ids_per_worker, geometry): it updated nothing");
    B200PS_REQUIRE(hdr[HDR_STUCK] == 0, B200PS_ERR_STATE, "a radix pass gave up waiting for an earlier tile: the plan is not valid");
    const uint32_t kept = hdr[HDR_KEPT], uniq = hdr[HDR_UNIQUE];
    if (n_kept) *n_kept = kept;
    if (n_unique) *n_unique = uniq;
    if (n_bad) *n_bad = hdr[HDR_BAD];
    const int fin = (key_passes(rows) - 1) & 1;  // the first pass wrote buffer 0, each later pass flips
    std::vector<uint32_t> tmp(kept > uniq + 1 ? kept : uniq + 1);
    if (h_sorted_ids && kept) {
        B200PS_CUDA(cudaMemcpyAsync(tmp.data(), ws.key[fin], kept * 4ull, cudaMemcpyDeviceToHost, st));
        B200PS_CUDA(cudaStreamSynchronize(st));
        for (uint32_t k = 0; k < kept; ++k) h_sorted_ids[k] = (int64_t)tmp[k];
    }
    if (h_perm && kept) B200PS_CUDA(cudaMemcpyAsync(h_perm, ws.val[fin], kept * 4ull, cudaMemcpyDeviceToHost, st));
    if (h_seg_id && kept) B200PS_CUDA(cudaMemcpyAsync(h_seg_id, ws.seg, kept * 4ull, cudaMemcpyDeviceToHost, st));
    if (h_seg_offset) B200PS_CUDA(cudaMemcpyAsync(h_seg_offset, ws.seg_off, (uniq + 1) * 4ull, cudaMemcpyDeviceToHost, st));
    B200PS_CUDA(cudaStreamSynchronize(st));
    if (h_unique_ids && uniq) {
        B200PS_CUDA(cudaMemcpyAsync(tmp.data(), ws.unique, uniq * 4ull, cudaMemcpyDeviceToHost, st));
        B200PS_CUDA(cudaStreamSynchronize(st));
        for (uint32_t k = 0; k < uniq; ++k) h_unique_ids[k] = (int64_t)tmp[k];
    }
    return B200PS_OK;
}

// ------------------------------------------------------------------------------------------------
// b200ps_sparse_push_sgd_noalloc: a whole sparse push (sort_segment + apply_sgd) for callers that
// must not allocate device memory.  The plan / sort scratch lives in the fixed array below instead
// of a caller-provided workspace.
//
// Who needs this: a PS or worker task whose share of HBM was handed out completely when the job
// was launched (table shards, gradient buffers, replicas -- the launcher sizes them to fill the
// GPU), so a later cudaMalloc may simply fail; and a step loop that keeps push, apply and pull in
// flight on several streams, where cudaMalloc / cudaFree would stall every one of them because they
// synchronise the whole device.  Such a caller passes only its inputs and the table it updates.
//
// The array is part of the library image: every GPU on which this process uses the library keeps
// B200PS_NOALLOC_SCRATCH_BYTES of it resident until the process exits, whether or not this entry
// point is ever called.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200PS_NOALLOC_SCRATCH_BYTES];

namespace {

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel can
// use, so the device address is looked up once per device (with that device current) and cached.
int noalloc_scratch_address(void** out) {
    static void* cached[64] = {nullptr};
    int dev = 0;
    B200PS_CUDA(cudaGetDevice(&dev));
    B200PS_REQUIRE(dev >= 0 && dev < 64, B200PS_ERR_RANGE, "device ordinal %d is outside 0..63", dev);
    void* p = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        B200PS_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        B200PS_REQUIRE(p && ((uintptr_t)p & 255) == 0, B200PS_ERR_CUDA, "static scratch has no usable device address");
        __atomic_store_n(&cached[dev], p, __ATOMIC_RELEASE);  // every thread resolves the same address
    }
    *out = p;
    return B200PS_OK;
}

}  // namespace

extern "C" int b200ps_sparse_push_sgd_noalloc(float* const* d_shard_base, const int64_t* const* d_ids,
                                              const float* const* d_grad_rows, int W, uint64_t ids_per_worker,
                                              const b200ps_table_geom_t* geom, float lr, void* stream) {
    B200PS_REQUIRE(W >= 1 && W <= B200PS_MAX_WORKERS, B200PS_ERR_RANGE, "W = %d is outside 1..%d", W, B200PS_MAX_WORKERS);
    const uint64_t n = (uint64_t)W * ids_per_worker;
    B200PS_REQUIRE(n >= 1 && n <= B200PS_NOALLOC_MAX_ENTRIES, B200PS_ERR_RANGE,
                   "push of %llu entries (W * ids_per_worker) exceeds the %llu the static scratch holds; "
                   "use b200ps_sparse_sort_segment + b200ps_sparse_apply_sgd with your own workspace",
                   (unsigned long long)n, (unsigned long long)B200PS_NOALLOC_MAX_ENTRIES);
    B200PS_REQUIRE(geom, B200PS_ERR_INVALID, "null table geometry");
    const size_t need = carve(nullptr, n, geom->dim > 0 ? geom->dim : 0).bytes;
    B200PS_REQUIRE(need <= (size_t)B200PS_NOALLOC_SCRATCH_BYTES, B200PS_ERR_RANGE,
                   "static scratch of %llu bytes cannot hold the %zu-byte plan of %llu entries",
                   (unsigned long long)B200PS_NOALLOC_SCRATCH_BYTES, need, (unsigned long long)n);
    void* scratch = nullptr;
    if (int rc = noalloc_scratch_address(&scratch)) return rc;
    // The array still holds the previous call's plan.  sort_segment zeroes the header (the only
    // part that must start at zero) on the stream and rewrites every other word it later reads.
    if (int rc = b200ps_sparse_sort_segment(d_ids, W, ids_per_worker, geom, scratch, B200PS_NOALLOC_SCRATCH_BYTES, stream))
        return rc;
    return b200ps_sparse_apply_sgd(d_shard_base, d_grad_rows, W, ids_per_worker, geom, lr, scratch, stream);
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU routing with posted NVLink writes.
//
// A random 256-byte read from peer memory exposes the full NVLink round trip (~3 us) per row and
// ran at ~90 GB/s; a write is posted and streams.  So rows never travel as peer reads:
//   pull  `serve`: the OWNER of a shard walks every worker's ids (streamed, local or peer), reads
//         the rows it owns from its local HBM and writes them into the requesting worker's output
//         buffer (a peer store when that worker lives on another GPU).  One launch serves all
//         workers; on one GPU it is simply the batched gather.
//   push  `route`: the WORKER's GPU walks its own ids and gradient rows and stores every row owned
//         by another GPU into that GPU's staging slot for this worker; the owner's apply then
//         reads local memory only.
// ------------------------------------------------------------------------------------------------
namespace b200ps {
namespace {

struct ServeArgs {
    float* out[B200PS_MAX_WORKERS];  // worker w's pulled rows [B, dim]
};

// Two phases per tile of 256 entries.  A: one THREAD per entry reads its id (coalesced), does the
// owner / worker arithmetic once and leaves a (source row, destination row) pair in shared memory.
// B: one LANE GROUP per row copies it.  (The first version did the arithmetic in every lane of the
// group: ncu showed 42 % of the issue slots busy with it.)
template <int LG>
__global__ void __launch_bounds__(256) sparse_serve_kernel(const __grid_constant__ ServeArgs sa,
                                                           const __grid_constant__ ShardBases tb,
                                                           const __grid_constant__ IdSources src,
                                                           const __grid_constant__ Geom g) {
    __shared__ const float4* s_src[256];
    __shared__ float4* s_dst[256];
    const int dim4 = g.dim / 4;
    for (uint32_t base = blockIdx.x * 256u; base < src.n_total; base += gridDim.x * 256u) {
        const uint32_t e = base + threadIdx.x;
        const float4* rd = nullptr;
        float4* wr = nullptr;
        if (e < src.n_total) {
            const uint32_t w = src.by_per_worker.div(e), pos = e - w * src.per_worker;
            const int64_t id = src.ids[w][pos];
            const bool in_range = id >= 0 && (uint64_t)id < g.rows;
            uint64_t lrow = 0;
            const int s = in_range ? owner_of(g, (uint64_t)id, &lrow) : 0;
            // an out-of-range id belongs to nobody: the GPU of shard 0 answers it with zeros
            if (gpu_of(g, s) == g.rank) {
                wr = reinterpret_cast<float4*>(sa.out[w] + (uint64_t)pos * (uint64_t)g.dim);
                if (in_range) rd = reinterpret_cast<const float4*>(tb.base[s] + lrow * (uint64_t)g.dim);
            }
        }
        s_src[threadIdx.x] = rd;
        s_dst[threadIdx.x] = wr;
        __syncthreads();
        const uint32_t left = src.n_total - base;
        copy_tile_rows<LG, 4>(s_src, s_dst, left < 256u ? (int)left : 256, dim4);
        __syncthreads();
    }
}

struct RouteArgs {
    float* dst[B200PS_MAX_MIRRORS + 1];  // per GPU: its staging slot for this worker (null for this GPU)
};

template <int LG>
__global__ void __launch_bounds__(256) sparse_route_kernel(const __grid_constant__ RouteArgs ra,
                                                           const float* __restrict__ grad_rows,
                                                           const int64_t* __restrict__ ids, uint32_t n,
                                                           const __grid_constant__ Geom g) {
    __shared__ const float4* s_src[256];
    __shared__ float4* s_dst[256];
    const int dim4 = g.dim / 4;
    for (uint32_t base = blockIdx.x * 256u; base < n; base += gridDim.x * 256u) {
        const uint32_t k = base + threadIdx.x;
        const float4* rd = nullptr;
        float4* wr = nullptr;
        if (k < n) {
            const int64_t id = ids[k];
            if (id >= 0 && (uint64_t)id < g.rows) {
                uint64_t lrow;
                const int o = gpu_of(g, owner_of(g, (uint64_t)id, &lrow));
                if (o != g.rank) {  // owned elsewhere: into that GPU's staging slot for this worker
                    rd = reinterpret_cast<const float4*>(grad_rows + (uint64_t)k * (uint64_t)g.dim);
                    wr = reinterpret_cast<float4*>(ra.dst[o] + (uint64_t)k * (uint64_t)g.dim);
                }
            }
        }
        s_src[threadIdx.x] = rd;
        s_dst[threadIdx.x] = wr;
        __syncthreads();
        const uint32_t left = n - base;
        copy_tile_rows<LG, 4>(s_src, s_dst, left < 256u ? (int)left : 256, dim4);
        __syncthreads();
    }
}

}  // namespace
}  // namespace b200ps

extern "C" int b200ps_sparse_serve(float* const* d_out, const float* const* d_shard_base, const int64_t* const* d_ids,
                                   int W, uint64_t ids_per_worker, const b200ps_table_geom_t* geom, void* stream) {
    Geom g;
    if (int rc = make_geom(geom, &g)) return rc;
    B200PS_REQUIRE(W >= 1 && W <= B200PS_MAX_WORKERS && d_out && d_ids && d_shard_base, B200PS_ERR_INVALID,
                   "bad worker count or null argument");
    const uint64_t n = (uint64_t)W * ids_per_worker;
    B200PS_REQUIRE(n < 0xFFFFFFFFull, B200PS_ERR_RANGE, "W * ids_per_worker must be below 2^32");
    if (n == 0) return B200PS_OK;
    ServeArgs sa{};
    IdSources src{};
    src.per_worker = (uint32_t)ids_per_worker;
    src.by_per_worker.set(src.per_worker);
    src.n_total = (uint32_t)n;
    for (int w = 0; w < W; ++w) {
        B200PS_REQUIRE(d_out[w] && ((uintptr_t)d_out[w] & 15) == 0 && d_ids[w], B200PS_ERR_INVALID,
                       "output / ids of worker %d: null or not 16-byte aligned", w);
        sa.out[w] = d_out[w];
        src.ids[w] = d_ids[w];
    }
    ShardBases tb{};
    for (int s = 0; s < g.nshards; ++s) {
        tb.base[s] = const_cast<float*>(d_shard_base[s]);
        if (s % g.world == g.rank)
            B200PS_REQUIRE(d_shard_base[s] && ((uintptr_t)d_shard_base[s] & 15) == 0, B200PS_ERR_INVALID,
                           "table shard %d must be a 16-byte aligned device pointer", s);
    }
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const int lg = lane_group(g.dim);
    cudaStream_t st = (cudaStream_t)stream;
    const uint64_t tiles = (n + 255) / 256;
    // one CTA per tile rather than a resident grid: measured a little faster, and slots free up all the
    // time for the sort of a table step that runs beside this kernel (table.cu)
#define LAUNCH_SERVE(LG) sparse_serve_kernel<LG><<<(unsigned)tiles, 256, 0, st>>>(sa, tb, src, g)
    switch (lg) {
        case 4: LAUNCH_SERVE(4); break;
        case 8: LAUNCH_SERVE(8); break;
        case 16: LAUNCH_SERVE(16); break;
        default: LAUNCH_SERVE(32); break;
    }
#undef LAUNCH_SERVE
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}

extern "C" int b200ps_sparse_route(const float* d_grad_rows, const int64_t* d_ids, uint64_t n,
                                   const b200ps_table_geom_t* geom, float* const* d_dst_by_gpu, void* stream) {
    Geom g;
    if (int rc = make_geom(geom, &g)) return rc;
    B200PS_REQUIRE(g.world <= B200PS_MAX_MIRRORS + 1, B200PS_ERR_RANGE, "at most %d GPUs", B200PS_MAX_MIRRORS + 1);
    B200PS_REQUIRE(n < 0xFFFFFFFFull, B200PS_ERR_RANGE, "n must be below 2^32");
    if (n == 0 || g.world == 1) return B200PS_OK;  // nothing lives elsewhere
    B200PS_REQUIRE(d_grad_rows && d_ids && d_dst_by_gpu && ((uintptr_t)d_grad_rows & 15) == 0, B200PS_ERR_INVALID,
                   "null or misaligned argument");
    RouteArgs ra{};
    for (int o = 0; o < g.world; ++o) {
        if (o == g.rank) continue;
        B200PS_REQUIRE(d_dst_by_gpu[o] && ((uintptr_t)d_dst_by_gpu[o] & 15) == 0, B200PS_ERR_INVALID,
                       "staging slot on GPU %d: null or not 16-byte aligned", o);
        ra.dst[o] = d_dst_by_gpu[o];
    }
    int sms = 0;
    B200PS_CUDA(sm_count(&sms));
    const int lg = lane_group(g.dim);
    cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH_ROUTE(LG) \
    sparse_route_kernel<LG><<<resident_grid(sparse_route_kernel<LG>, 256, (n + 255) / 256), 256, 0, st>>>(ra, d_grad_rows, d_ids, (uint32_t)n, g)
    switch (lg) {
        case 4: LAUNCH_ROUTE(4); break;
        case 8: LAUNCH_ROUTE(8); break;
        case 16: LAUNCH_ROUTE(16); break;
        default: LAUNCH_ROUTE(32); break;
    }
#undef LAUNCH_ROUTE
    B200PS_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return B200PS_OK;
}
