// sort.cuh -- hand-written LSD radix sort (8-bit digits) for the HLBVH build, plus a device
// exclusive scan.  Replaces the reference's one-pass atomic counting sort on the top 21 Morton
// bits (accelerator/hlbvh.cu:56-82, host scan :256-259), whose order inside a bucket is the
// atomic arrival order; this sort is stable and therefore deterministic.
//
// Per pass: k_sort_upsweep (per-tile digit histogram) -> exclusive scan of the digit-major
// [256][tiles] table -> k_sort_downsweep (stable warp-level multi-split ranking with
// __match_any_sync, scatter).  Keys are uint64; an optional uint32 payload rides along.
#pragma once

#include <algorithm>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200pt {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // 4096 keys per CTA
constexpr int SORT_WARPS = SORT_THREADS / 32;

// ---------------------------------------------------------------------------- exclusive scan
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread; returns the exclusive prefix and the total
template <int THREADS>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t &total, uint32_t *smem /*[THREADS/32]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t incl = warp_incl_scan(v);
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < THREADS / 32) ? smem[lane] : 0;
        const uint32_t wi = warp_incl_scan(w);
        if (lane < THREADS / 32) smem[lane] = wi - w;
        if (lane == THREADS / 32 - 1) smem[THREADS / 32] = wi;
    }
    __syncthreads();
    total = smem[THREADS / 32];
    const uint32_t r = smem[warp] + incl - v;
    __syncthreads();
    return r;
}

// phase 1: per-tile sums
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const uint32_t *__restrict__ in, int64_t n,
                                                              uint32_t *__restrict__ tile_sums) {
    __shared__ uint32_t sm[SCAN_THREADS / 32 + 1];
    const int64_t base = int64_t(blockIdx.x) * SCAN_TILE;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        const int64_t k = base + i * SCAN_THREADS + threadIdx.x;
        if (k < n) s += in[k];
    }
    uint32_t total;
    block_excl_scan<SCAN_THREADS>(s, total, sm);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// phase 2: one CTA scans the tile sums in place (exclusive), chunk by chunk
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_spine(uint32_t *__restrict__ tile_sums, int n_tiles) {
    __shared__ uint32_t sm[SCAN_THREADS / 32 + 1];
    uint32_t carry = 0;
    for (int base = 0; base < n_tiles; base += SCAN_THREADS) {
        const int k = base + threadIdx.x;
        const uint32_t v = (k < n_tiles) ? tile_sums[k] : 0;
        uint32_t total;
        const uint32_t ex = block_excl_scan<SCAN_THREADS>(v, total, sm);
        if (k < n_tiles) tile_sums[k] = carry + ex;
        carry += total;
    }
}

// phase 3: per-tile exclusive scan + tile offset.  Thread t owns SCAN_ITEMS consecutive items.
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const uint32_t *__restrict__ in, uint32_t *__restrict__ out,
                                                             int64_t n, const uint32_t *__restrict__ tile_sums) {
    __shared__ uint32_t sm[SCAN_THREADS / 32 + 1];
    const int64_t base = int64_t(blockIdx.x) * SCAN_TILE + int64_t(threadIdx.x) * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        v[i] = (base + i < n) ? in[base + i] : 0;
        s += v[i];
    }
    uint32_t total;
    uint32_t ex = block_excl_scan<SCAN_THREADS>(s, total, sm) + tile_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        if (base + i < n) out[base + i] = ex;
        ex += v[i];
    }
}

// exclusive scan of n uint32 (in place allowed); tmp needs ceil(n / SCAN_TILE) + 1 words
inline void exclusive_scan_u32(const uint32_t *in, uint32_t *out, int64_t n, uint32_t *tmp, cudaStream_t st) {
    if (n <= 0) return;
    const int tiles = int((n + SCAN_TILE - 1) / SCAN_TILE);
    k_scan_reduce<<<tiles, SCAN_THREADS, 0, st>>>(in, n, tmp);
    k_scan_spine<<<1, SCAN_THREADS, 0, st>>>(tmp, tiles);
    k_scan_apply<<<tiles, SCAN_THREADS, 0, st>>>(in, out, n, tmp);
}

// ------------------------------------------------------------------------------- radix sort
__global__ void __launch_bounds__(SORT_THREADS) k_sort_upsweep(const uint64_t *__restrict__ keys, int64_t n, int shift,
                                                               uint32_t *__restrict__ tile_hist, int n_tiles) {
    __shared__ uint32_t hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = int64_t(blockIdx.x) * SORT_TILE;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = base + i * SORT_THREADS + threadIdx.x;
        if (k < n) atomicAdd(&hist[(keys[k] >> shift) & 255u], 1u);
    }
    __syncthreads();
    tile_hist[int64_t(threadIdx.x) * n_tiles + blockIdx.x] = hist[threadIdx.x];
}

template <bool HAS_VALUES>
__global__ void __launch_bounds__(SORT_THREADS) k_sort_downsweep(const uint64_t *__restrict__ keys_in, uint64_t *__restrict__ keys_out,
                                                                 const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out,
                                                                 int64_t n, int shift, const uint32_t *__restrict__ tile_offsets,
                                                                 int n_tiles) {
    __shared__ uint32_t cnt[SORT_WARPS][256]; // per-warp digit counts -> per-warp exclusive offsets
    __shared__ uint32_t gbase[256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < SORT_WARPS * 256; i += SORT_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();

    // warp w owns keys [base + w*512, +512), item i of lane l at i*32 + l: (item, lane) order is
    // the input order, so ranking in that order keeps the sort stable.
    const int64_t wbase = int64_t(blockIdx.x) * SORT_TILE + warp * (32 * SORT_ITEMS);
    uint64_t key[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = wbase + i * 32 + lane;
        const bool valid = k < n;
        key[i] = valid ? keys_in[k] : ~0ull;
        const uint32_t d = valid ? uint32_t((key[i] >> shift) & 255u) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (lane == leader && valid) {
            prev = cnt[warp][d];
            cnt[warp][d] = prev + __popc(peers);
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[i] = prev + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    { // digit d = threadIdx.x: exclusive offsets over warps, then the tile's global base
        const int d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            const uint32_t c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
        gbase[d] = tile_offsets[int64_t(d) * n_tiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = wbase + i * 32 + lane;
        if (k < n) {
            const uint32_t d = uint32_t((key[i] >> shift) & 255u);
            const uint32_t pos = gbase[d] + cnt[warp][d] + rank[i];
            keys_out[pos] = key[i];
            if (HAS_VALUES) vals_out[pos] = vals_in[k];
        }
    }
}

// lanes of the warp holding the same 8-bit digit (256 marks an invalid lane: it matches only other invalid lanes): eight ballots
// and a few logic operations instead of MATCH.ANY, whose issue rate is a fraction of theirs
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
    uint32_t peers = __ballot_sync(0xffffffffu, d > 255u);
    peers = d > 255u ? peers : ~peers;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const uint32_t m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? m : ~m;
    }
    return peers;
}

// ---- one-sweep form (key-only): the digit histograms of ALL passes come from one read of the keys, and a pass is ONE kernel
// that reads and writes each key once -- a tile ranks its keys, publishes its digit counts and finds its global offsets by
// looking back over the tiles before it (decoupled look-back) instead of an upsweep, a three-kernel scan and a downsweep.
// Tiles take their numbers from an atomic counter in the order they start, so every tile a tile waits for is already running.
constexpr int SORT_MAX_PASSES = 8;
#ifndef SORT_BALLOT_MATCH
#define SORT_BALLOT_MATCH 1
#endif
constexpr unsigned long long SORT_FLAG_AGG = 1ull << 62, SORT_FLAG_INCL = 2ull << 62, SORT_VALUE_MASK = (1ull << 62) - 1;

__global__ void __launch_bounds__(SORT_THREADS) k_sort_histograms(const uint64_t *__restrict__ keys, int64_t n, int bit_lo, int passes,
                                                                  uint32_t *__restrict__ hist /* [passes][256] */) {
    __shared__ uint32_t sh[SORT_MAX_PASSES * 256];
    for (int i = threadIdx.x; i < passes * 256; i += SORT_THREADS) sh[i] = 0;
    __syncthreads();
    for (int64_t k = int64_t(blockIdx.x) * SORT_THREADS + threadIdx.x; k < n; k += int64_t(gridDim.x) * SORT_THREADS) {
        const uint64_t key = keys[k] >> bit_lo;
        for (int p = 0; p < passes; p++) atomicAdd(&sh[p * 256 + int((key >> (8 * p)) & 255u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += SORT_THREADS)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// hist[p][d] -> exclusive prefix over d, one CTA of 256 threads per pass
__global__ void __launch_bounds__(256) k_sort_digit_offsets(uint32_t *__restrict__ hist) {
    __shared__ uint32_t sh[256];
    uint32_t *h = hist + blockIdx.x * 256;
    const uint32_t v = h[threadIdx.x];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
        const uint32_t t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0u;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    h[threadIdx.x] = sh[threadIdx.x] - v;
}

__global__ void __launch_bounds__(SORT_THREADS) k_sort_onesweep(const uint64_t *__restrict__ keys_in, uint64_t *__restrict__ keys_out, int64_t n,
                                                                int shift, const uint32_t *__restrict__ digit_offsets /* [256] */,
                                                                unsigned long long *tile_state /* [tiles][256], zeroed */,
                                                                unsigned int *tile_counter /* zeroed */) {
    __shared__ uint32_t cnt[SORT_WARPS][256]; // per-warp digit counts -> per-warp exclusive offsets
    __shared__ uint32_t gbase[256];           // global position of the tile's first key of digit d, minus dstart[d]
    __shared__ uint32_t dstart[256];          // position of that key in the tile's own sorted order
    __shared__ uint32_t warp_tot[SORT_WARPS];
    __shared__ uint64_t staged[SORT_TILE];    // the tile's keys in sorted order: the global writes become contiguous runs
    __shared__ unsigned int s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1u);
    for (int i = threadIdx.x; i < SORT_WARPS * 256; i += SORT_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const unsigned int tile = s_tile;

    // warp w owns keys [base + w*512, +512), item i of lane l at i*32 + l: (item, lane) order is
    // the input order, so ranking in that order keeps the sort stable.
    const int64_t wbase = int64_t(tile) * SORT_TILE + warp * (32 * SORT_ITEMS);
    uint64_t key[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = wbase + i * 32 + lane;
        key[i] = k < n ? keys_in[k] : ~0ull;
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = wbase + i * 32 + lane;
        const bool valid = k < n;
        const uint32_t d = valid ? uint32_t((key[i] >> shift) & 255u) : 256u;
        const uint32_t peers = SORT_BALLOT_MATCH ? match_digit(d) : __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (lane == leader && valid) {
            prev = cnt[warp][d];
            cnt[warp][d] = prev + __popc(peers);
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[i] = prev + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    { // digit d = threadIdx.x: exclusive offsets over warps, the tile's count, then the look-back for its global base
        const int d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            const uint32_t c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
        volatile unsigned long long *state = tile_state;
        unsigned long long excl = 0;
        if (tile == 0) {
            state[d] = SORT_FLAG_INCL | run;
        } else {
            state[size_t(tile) * 256 + d] = SORT_FLAG_AGG | run;
            for (long long t = (long long)tile - 1; t >= 0; t--) {
                unsigned long long v;
                do {
                    v = state[size_t(t) * 256 + d];
                } while ((v & ~SORT_VALUE_MASK) == 0ull);
                excl += v & SORT_VALUE_MASK;
                if (v & SORT_FLAG_INCL) break;
            }
            state[size_t(tile) * 256 + d] = SORT_FLAG_INCL | (excl + run);
        }
        // where digit d starts inside the tile: block-wide exclusive scan of the 256 tile counts
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) before += w < warp ? warp_tot[w] : 0u;
        const uint32_t start = before + incl - run;
        dstart[d] = start;
        gbase[d] = digit_offsets[d] + uint32_t(excl) - start;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t k = wbase + i * 32 + lane;
        if (k < n) {
            const uint32_t d = uint32_t((key[i] >> shift) & 255u);
            staged[dstart[d] + cnt[warp][d] + rank[i]] = key[i];
        }
    }
    __syncthreads();
    const int64_t tile_first = int64_t(tile) * SORT_TILE;
    const int tile_n = int(n - tile_first < SORT_TILE ? n - tile_first : SORT_TILE);
    for (int j = threadIdx.x; j < tile_n; j += SORT_THREADS) {
        const uint64_t kv = staged[j];
        keys_out[gbase[uint32_t((kv >> shift) & 255u)] + j] = kv;
    }
}

struct SortBuffers {
    uint64_t *keys[2] = {nullptr, nullptr};
    uint32_t *vals[2] = {nullptr, nullptr}; // null => key-only sort
    uint32_t *tile_hist = nullptr;          // [256 * n_tiles]
    uint32_t *scan_tmp = nullptr;           // [ceil(256*n_tiles / SCAN_TILE) + 1]
    // one-sweep form (key-only sorts): sort_onesweep_bytes(n, passes) bytes, 8-byte aligned; null => the three-kernel form
    void *onesweep = nullptr;
};

inline int64_t sort_tiles(int64_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }
inline int64_t sort_scan_tmp_words(int64_t n) { return (256 * sort_tiles(n) + SCAN_TILE - 1) / SCAN_TILE + 1; }

// layout of the one-sweep scratch: [passes][tiles][256] u64 tile states, [passes][256] u32 histograms, [passes] u32 tile counters
inline size_t sort_onesweep_bytes(int64_t n, int passes) {
    return size_t(passes) * size_t(sort_tiles(n)) * 256 * 8 + size_t(passes) * 256 * 4 + size_t(SORT_MAX_PASSES) * 4;
}

// Sorts bits [bit_lo, bit_hi) (a multiple of 8 wide).  Returns the index (0/1) of the buffer
// holding the result.
inline int radix_sort_u64(SortBuffers &b, int64_t n, int bit_lo, int bit_hi, cudaStream_t st, int *passes_out, int sm_count = 148) {
    const int n_tiles = int(sort_tiles(n));
    int cur = 0, passes = 0;
    const int want = (bit_hi - bit_lo) / 8;
    if (b.onesweep && !b.vals[0] && want >= 1 && want <= SORT_MAX_PASSES) {
        unsigned long long *state = (unsigned long long *)b.onesweep;
        uint32_t *hist = (uint32_t *)(state + size_t(want) * n_tiles * 256);
        unsigned int *counters = hist + want * 256;
        cudaMemsetAsync(b.onesweep, 0, sort_onesweep_bytes(n, want), st);
        const int hgrid = int(std::min<int64_t>((n + SORT_THREADS - 1) / SORT_THREADS, int64_t(sm_count) * 8));
        k_sort_histograms<<<hgrid, SORT_THREADS, 0, st>>>(b.keys[0], n, bit_lo, want, hist);
        k_sort_digit_offsets<<<want, 256, 0, st>>>(hist);
        for (int p = 0; p < want; p++) {
            k_sort_onesweep<<<n_tiles, SORT_THREADS, 0, st>>>(b.keys[cur], b.keys[cur ^ 1], n, bit_lo + 8 * p, hist + p * 256,
                                                              state + size_t(p) * n_tiles * 256, counters + p);
            cur ^= 1;
        }
        if (passes_out) *passes_out = want;
        return cur;
    }
    for (int shift = bit_lo; shift < bit_hi; shift += 8, passes++) {
        k_sort_upsweep<<<n_tiles, SORT_THREADS, 0, st>>>(b.keys[cur], n, shift, b.tile_hist, n_tiles);
        exclusive_scan_u32(b.tile_hist, b.tile_hist, int64_t(256) * n_tiles, b.scan_tmp, st);
        if (b.vals[0])
            k_sort_downsweep<true><<<n_tiles, SORT_THREADS, 0, st>>>(b.keys[cur], b.keys[cur ^ 1], b.vals[cur], b.vals[cur ^ 1],
                                                                     n, shift, b.tile_hist, n_tiles);
        else
            k_sort_downsweep<false><<<n_tiles, SORT_THREADS, 0, st>>>(b.keys[cur], b.keys[cur ^ 1], nullptr, nullptr, n, shift,
                                                                      b.tile_hist, n_tiles);
        cur ^= 1;
    }
    if (passes_out) *passes_out = passes;
    return cur;
}

} // namespace b200pt
