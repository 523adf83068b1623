// reduce_kernels.cuh -- everything after the hash-combine, on DISTINCT keys only:
//   K3  compact the global table into entry arrays, partition id per key (exact Lua
//       5.2 double FNV, partitionfn.lua:8-16) and the 96-bit sort key (partition,
//       first 7 key bytes big-endian, min(len, 8))
//   K5  LSD radix sort on the bits that actually vary (replaces utils.keys_sorted,
//       utils.lua:123-128, and the heap merge order, utils.lua:206-271 + heap.lua)
//       + resolution of keys that agree in their first 7 bytes
//   K6  segmented reduce: equal neighbours (same key from several sources) are summed
//       (merge_min_keys utils.lua:232-247 + reducefn.lua:1-5)
//   K7  rows: partition, count, key offsets, packed key bytes, partition starts
#pragma once
#include "common.cuh"

namespace mr {

constexpr int RADIX_BITS = 8;
constexpr int RADIX_BINS = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 12;          // 96 bits / 8
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // 4096
constexpr int SCAN_TILE = 4096;
constexpr uint32_t TIE_SMALL_MAX = 1024; // tie entries resolved inside one CTA

struct DigitRun { uint8_t shift, width, dst, pad; }; // shift 0..63 -> lo, 64..95 -> hi

struct Ctl {
    unsigned long long n;            // entries (distinct keys before the merge)
    unsigned long long or_lo, and_lo;
    unsigned int or_hi, and_hi;
    unsigned int npasses;
    unsigned int err;
    unsigned int n_tie;              // entries whose order needs bytes beyond the first 7
    unsigned int tie_overflow;       // n_tie > TIE_SMALL_MAX: host must run the general path
    unsigned long long n_out;        // rows after the segmented reduce
    unsigned long long key_bytes;    // sum of key lengths over rows
    unsigned long long tokens;
    unsigned int nruns[MAX_PASSES];
    DigitRun runs[MAX_PASSES][8];
    unsigned int tickets[MAX_PASSES];
    unsigned int ghist[MAX_PASSES][RADIX_BINS];
};

// ------------------------------------------------------------------ K3
struct EntryArrays {
    unsigned long long *k0, *k1; // key words (common.cuh)
    unsigned long long *cnt;
    uint32_t *part;
    __device__ __forceinline__ Key key(uint64_t i) const { Key k; k.k0 = k0[i]; k.k1 = k1[i]; return k; }
};

__device__ __forceinline__ void orand_commit(Ctl *ctl, uint32_t hi_or, uint32_t hi_and,
                                             uint64_t lo_or, uint64_t lo_and, bool any)
{
    // warp-reduce then one atomic per warp
    for (int d = 16; d; d >>= 1) {
        uint32_t o_hi_or = __shfl_xor_sync(0xFFFFFFFFu, hi_or, d);
        uint32_t o_hi_and = __shfl_xor_sync(0xFFFFFFFFu, hi_and, d);
        uint64_t o_lo_or = __shfl_xor_sync(0xFFFFFFFFu, lo_or, d);
        uint64_t o_lo_and = __shfl_xor_sync(0xFFFFFFFFu, lo_and, d);
        hi_or |= o_hi_or; hi_and &= o_hi_and; lo_or |= o_lo_or; lo_and &= o_lo_and;
    }
    uint32_t anym = __ballot_sync(0xFFFFFFFFu, any);
    if ((threadIdx.x & 31) == 0 && anym) {
        atomicOr(&ctl->or_hi, hi_or);
        atomicAnd(&ctl->and_hi, hi_and);
        atomicOr(&ctl->or_lo, (unsigned long long)lo_or);
        atomicAnd(&ctl->and_lo, (unsigned long long)lo_and);
    }
}

// table slots -> dense entries.  Order of entries is arbitrary.  One atomic per 1024-slot
// tile reserves the output range (a per-warp atomic on one address costs ~1.4 ns each).
constexpr int COMPACT_ITEMS = 4;
__global__ void __launch_bounds__(256) compact_kernel(const GEntry *gt, uint64_t nslots, EntryArrays ent,
                                                      uint64_t max_keys, Ctl *ctl)
{
    __shared__ unsigned int wsum[8];
    __shared__ unsigned long long s_base;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t tile_slots = 256 * COMPACT_ITEMS;
    const uint64_t ntiles = (nslots + tile_slots - 1) / tile_slots;
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        Key key[COMPACT_ITEMS];
        uint32_t nocc = 0;
#pragma unroll
        for (int i = 0; i < COMPACT_ITEMS; ++i) {
            uint64_t slot = tile * tile_slots + (uint64_t)i * 256 + threadIdx.x;
            key[i].k0 = 0; key[i].k1 = 0;
            if (slot < nslots) {
                ulonglong2 v = __ldcg((const ulonglong2 *)&gt[slot]);
                key[i].k0 = v.x; key[i].k1 = v.y;
            }
            nocc += key[i].k1 != 0;
        }
        uint32_t incl = nocc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        uint32_t woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { uint32_t c = wsum[w]; if (w < (int)warp) woff += c; total += c; }
        if (threadIdx.x == 0 && total) s_base = atomicAdd(&ctl->n, (unsigned long long)total);
        __syncthreads();
        if (total) {
            uint64_t idx = s_base + woff + incl - nocc;
#pragma unroll
            for (int i = 0; i < COMPACT_ITEMS; ++i) {
                if (key[i].k1 == 0) continue;
                uint64_t slot = tile * tile_slots + (uint64_t)i * 256 + threadIdx.x;
                if (idx < max_keys) {
                    ent.k0[idx] = key[i].k0; ent.k1[idx] = key[i].k1; ent.cnt[idx] = __ldcg(&gt[slot].cnt);
                } else {
                    atomicOr(&ctl->err, ERR_KEYS_FULL);
                }
                ++idx;
            }
        }
        __syncthreads();
    }
}

// entries already dense (after a shuffle import): partition + sort key only
__global__ void __launch_bounds__(256) partition_entries_kernel(
    const uint8_t *bytes, uint32_t num_reducers, EntryArrays ent, uint32_t *sk_hi,
    unsigned long long *sk_lo, uint32_t *val, uint64_t max_keys, Ctl *ctl)
{
    uint32_t hi_or = 0, hi_and = ~0u;
    uint64_t lo_or = 0, lo_and = ~0ull;
    bool any = false;
    const uint64_t n = ctl->n < max_keys ? ctl->n : max_keys;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        Key key = ent.key(idx);
        uint32_t part = partition_of_key(key, bytes, num_reducers);
        uint64_t chunk = sort_chunk(key, bytes, 0);
        ent.part[idx] = part;
        sk_hi[idx] = part; sk_lo[idx] = chunk; val[idx] = (uint32_t)idx;
        hi_or |= part; hi_and &= part; lo_or |= chunk; lo_and &= chunk;
        any = true;
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any);
}

// ------------------------------------------------------------------ K5 radix sort
// One thread: split the varying bits (or ^ and) into digits of up to 8 bits, least
// significant first; each digit is up to 8 runs of consecutive key bits.
__global__ void sort_setup_kernel(Ctl *ctl, unsigned long long max_keys)
{
    if (threadIdx.x || blockIdx.x) return;
    if (ctl->n > max_keys) { ctl->n = max_keys; ctl->err |= ERR_KEYS_FULL; } // reported by the host
    uint64_t vlo = ctl->or_lo ^ ctl->and_lo;
    uint32_t vhi = ctl->or_hi ^ ctl->and_hi;
    if (ctl->n < 2) { vlo = 0; vhi = 0; }
    unsigned int pass = 0, used = 0, nr = 0;
    int bit = 0;
    while (bit < 96) {
        bool set = bit < 64 ? ((vlo >> bit) & 1) : ((vhi >> (bit - 64)) & 1);
        if (!set) { ++bit; continue; }
        int start = bit, w = 0;
        int limit = start < 64 ? 64 : 96; // runs do not straddle the lo/hi boundary
        while (bit < limit && used + w < RADIX_BITS &&
               (bit < 64 ? ((vlo >> bit) & 1) : ((vhi >> (bit - 64)) & 1))) { ++bit; ++w; }
        DigitRun r; r.shift = (uint8_t)start; r.width = (uint8_t)w; r.dst = (uint8_t)used; r.pad = 0;
        ctl->runs[pass][nr++] = r;
        used += w;
        if (used == RADIX_BITS) { ctl->nruns[pass] = nr; ++pass; used = 0; nr = 0; }
    }
    if (used) { ctl->nruns[pass] = nr; ++pass; }
    ctl->npasses = pass;
}

__device__ __forceinline__ uint32_t extract_digit(const DigitRun *runs, uint32_t nruns, uint32_t hi,
                                                  uint64_t lo)
{
    uint32_t d = 0;
    for (uint32_t i = 0; i < nruns; ++i) {
        DigitRun r = runs[i];
        uint32_t src = r.shift < 64 ? (uint32_t)(lo >> r.shift) : (hi >> (r.shift - 64));
        d |= (src & ((1u << r.width) - 1u)) << r.dst;
    }
    return d;
}

// digit histograms of every pass in one read of the keys
__global__ void __launch_bounds__(256) radix_hist_kernel(const uint32_t *sk_hi,
                                                         const unsigned long long *sk_lo, Ctl *ctl)
{
    __shared__ unsigned int h[MAX_PASSES][RADIX_BINS];
    __shared__ DigitRun runs[MAX_PASSES][8];
    __shared__ unsigned int nruns[MAX_PASSES];
    const uint32_t np = ctl->npasses;
    const uint64_t n = ctl->n;
    if ((uint64_t)blockIdx.x * blockDim.x >= n || np == 0) return;
    for (uint32_t i = threadIdx.x; i < MAX_PASSES * RADIX_BINS; i += blockDim.x) (&h[0][0])[i] = 0;
    if (threadIdx.x < MAX_PASSES * 8) (&runs[0][0])[threadIdx.x] = (&ctl->runs[0][0])[threadIdx.x];
    if (threadIdx.x < MAX_PASSES) nruns[threadIdx.x] = ctl->nruns[threadIdx.x];
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t hi = sk_hi[i];
        uint64_t lo = sk_lo[i];
        for (uint32_t p = 0; p < np; ++p) atomicAdd(&h[p][extract_digit(runs[p], nruns[p], hi, lo)], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < np * RADIX_BINS; i += blockDim.x) {
        unsigned int c = (&h[0][0])[i];
        if (c) atomicAdd(&(&ctl->ghist[0][0])[i], c);
    }
}

// tile status word: bits 30..31 = state (0 empty, 1 aggregate, 2 inclusive prefix)
constexpr uint32_t ST_AGG = 1u << 30, ST_INC = 2u << 30, ST_VAL = (1u << 30) - 1;

struct SortBufs {
    uint32_t *hi[2];
    unsigned long long *lo[2];
    uint32_t *val[2];
};

// One LSD pass (stable): tile-local ranking by warp-level digit matching, tile offsets
// by decoupled look-back over per-tile digit counts.
__global__ void __launch_bounds__(SORT_THREADS) radix_pass_kernel(SortBufs bufs, uint32_t pass,
                                                                  uint32_t *status, uint32_t max_tiles,
                                                                  Ctl *ctl)
{
    if (pass >= ctl->npasses) return;
    const uint64_t n = ctl->n;
    __shared__ uint32_t s_tile;
    __shared__ unsigned int whist[SORT_THREADS / 32][RADIX_BINS];
    __shared__ unsigned int bin_base[RADIX_BINS];
    __shared__ unsigned int scan_tmp[SORT_THREADS / 32];
    __shared__ DigitRun runs[8];
    __shared__ unsigned int s_nruns;
    if (threadIdx.x == 0) s_tile = atomicAdd(&ctl->tickets[pass], 1u);
    if (threadIdx.x < 8) runs[threadIdx.x] = ctl->runs[pass][threadIdx.x];
    if (threadIdx.x == 8) s_nruns = ctl->nruns[pass];
    for (uint32_t i = threadIdx.x; i < (SORT_THREADS / 32) * RADIX_BINS; i += SORT_THREADS)
        (&whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    if ((uint64_t)tile * SORT_TILE >= n) return;
    const uint32_t src = pass & 1, dst = src ^ 1;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt = lanemask_lt();
    const uint32_t nruns = s_nruns;

    uint32_t k_hi[SORT_ITEMS], k_val[SORT_ITEMS];
    uint64_t k_lo[SORT_ITEMS];
    uint16_t k_rank[SORT_ITEMS], k_dig[SORT_ITEMS];
    const uint64_t wbase = (uint64_t)tile * SORT_TILE + (uint64_t)warp * (32 * SORT_ITEMS);
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        if (valid) { k_hi[i] = bufs.hi[src][idx]; k_lo[i] = bufs.lo[src][idx]; k_val[i] = bufs.val[src][idx]; }
        else { k_hi[i] = 0; k_lo[i] = 0; k_val[i] = 0; }
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        uint32_t act = __ballot_sync(0xFFFFFFFFu, valid);
        uint32_t d = 0, r = 0;
        if (valid) {
            d = extract_digit(runs, nruns, k_hi[i], k_lo[i]);
            uint32_t peers = __match_any_sync(act, d);
            uint32_t leader = __ffs((int)peers) - 1;
            uint32_t old = 0;
            if (lane == leader) { old = whist[warp][d]; whist[warp][d] = old + __popc(peers); }
            old = __shfl_sync(peers, old, leader);
            r = old + __popc(peers & lt);
        }
        k_dig[i] = (uint16_t)d; k_rank[i] = (uint16_t)r;
        __syncwarp();
    }
    __syncthreads();
    // thread b owns bin b: per-warp exclusive offsets, tile count, look-back
    {
        const uint32_t b = threadIdx.x;
        unsigned int run = 0;
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; ++w) { unsigned int c = whist[w][b]; whist[w][b] = run; run += c; }
        const unsigned int tile_count = run;
        volatile uint32_t *st = status + ((uint64_t)pass * max_tiles) * RADIX_BINS;
        if (tile == 0) st[(uint64_t)tile * RADIX_BINS + b] = ST_INC | tile_count;
        else st[(uint64_t)tile * RADIX_BINS + b] = ST_AGG | tile_count;
        __threadfence();
        unsigned int prefix = 0;
        if (tile > 0) {
            // look back 8 tiles at a time: the 8 loads are independent, so a long chain of
            // not-yet-inclusive predecessors costs one L2 round trip per 8 tiles, not per tile
            int t = (int)tile - 1;
            bool done = false;
            while (!done) {
                uint32_t sv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) sv[k] = t - k >= 0 ? st[(uint64_t)(t - k) * RADIX_BINS + b] : (2u << 30);
                int k = 0;
#pragma unroll
                for (; k < 8; ++k) {
                    if ((sv[k] >> 30) == 0) break;           // not published yet: re-read from here
                    prefix += sv[k] & ST_VAL;
                    if ((sv[k] >> 30) == 2) { done = true; break; }
                }
                t -= k;
            }
            st[(uint64_t)tile * RADIX_BINS + b] = ST_INC | (prefix + tile_count);
        }
        // exclusive scan of the global histogram = start of every bin
        unsigned int g = ctl->ghist[pass][b];
        unsigned int incl = g;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) scan_tmp[warp] = incl;
        __syncthreads();
        unsigned int wsum = 0;
        for (uint32_t w = 0; w < warp; ++w) wsum += scan_tmp[w];
        bin_base[b] = wsum + incl - g + prefix;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        if (idx < n) {
            uint32_t d = k_dig[i];
            uint64_t o = (uint64_t)bin_base[d] + whist[warp][d] + k_rank[i];
            bufs.hi[dst][o] = k_hi[i]; bufs.lo[dst][o] = k_lo[i]; bufs.val[dst][o] = k_val[i];
        }
    }
}

// ------------------------------------------------------------------ ties beyond 7 bytes
struct TieItem { uint32_t pos, val; };

__device__ inline bool tie_less(const TieItem &a, const TieItem &b, const EntryArrays &ent,
                                const uint8_t *bytes)
{
    if (a.pos == 0xFFFFFFFFu || b.pos == 0xFFFFFFFFu) return a.pos != 0xFFFFFFFFu && b.pos == 0xFFFFFFFFu;
    uint32_t pa = ent.part[a.val], pb = ent.part[b.val];
    if (pa != pb) return pa < pb;
    const Key ka = ent.key(a.val), kb = ent.key(b.val);
    // round by round, 7 key bytes at a time (round 0 is what the radix sort already used)
    for (uint32_t r = 0;; ++r) {
        uint64_t ca = sort_chunk(ka, bytes, r), cb = sort_chunk(kb, bytes, r);
        if (ca != cb) return ca < cb;
        if ((ca & 0xFF) < 8) break; // both keys end here: equal
    }
    return a.val < b.val;
}

// Marks sorted positions whose (partition, first 7 bytes) equals a neighbour's while
// more key bytes remain, and appends them (unordered) to the tie list.
__global__ void __launch_bounds__(256) tie_mark_kernel(SortBufs bufs, TieItem *ties, uint32_t tie_cap,
                                                       Ctl *ctl)
{
    const uint64_t n = ctl->n;
    const uint32_t fin = ctl->npasses & 1;
    const uint32_t *hi = bufs.hi[fin];
    const unsigned long long *lo = bufs.lo[fin];
    const uint32_t lane = threadIdx.x & 31, lt = lanemask_lt();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t rounds = (n + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; ++r) {
        uint64_t i = r * stride + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        bool tie = false;
        if (i < n) {
            uint64_t l = lo[i];
            if ((l & 0xFF) == 8) {
                uint32_t h = hi[i];
                tie = (i > 0 && hi[i - 1] == h && lo[i - 1] == l) ||
                      (i + 1 < n && hi[i + 1] == h && lo[i + 1] == l);
            }
        }
        uint32_t m = __ballot_sync(0xFFFFFFFFu, tie);
        if (!m) continue;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&ctl->n_tie, (unsigned int)__popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (tie) {
            uint32_t k = base + __popc(m & lt);
            if (k < tie_cap) { ties[k].pos = (uint32_t)i; ties[k].val = bufs.val[fin][i]; }
        }
    }
}

// Up to TIE_SMALL_MAX ties: one CTA copies their keys into shared memory, orders the positions
// and the entries (full key compare) with bitonic networks and writes the entries back.
__global__ void __launch_bounds__(1024) tie_fix_small_kernel(SortBufs bufs, const TieItem *ties,
                                                             EntryArrays ent, const uint8_t *bytes,
                                                             Ctl *ctl)
{
    const uint32_t nt = ctl->n_tie;
    if (nt == 0) return;
    if (nt > TIE_SMALL_MAX) { if (threadIdx.x == 0) ctl->tie_overflow = 1; return; }
    __shared__ Key skey[TIE_SMALL_MAX];
    __shared__ uint32_t spart[TIE_SMALL_MAX], sval[TIE_SMALL_MAX], posv[TIE_SMALL_MAX];
    __shared__ uint16_t order[TIE_SMALL_MAX];
    uint32_t np2 = 1;
    while (np2 < nt) np2 <<= 1;
    for (uint32_t i = threadIdx.x; i < np2; i += blockDim.x) {
        order[i] = (uint16_t)i;
        posv[i] = 0xFFFFFFFFu;
        if (i < nt) {
            TieItem t = ties[i];
            posv[i] = t.pos; sval[i] = t.val; spart[i] = ent.part[t.val]; skey[i] = ent.key(t.val);
        }
    }
    __syncthreads();
    auto less = [&](uint32_t a, uint32_t b) -> bool { // a, b index the unsorted copies; >= nt is padding
        if (a >= nt || b >= nt) return a < nt && b >= nt;
        if (spart[a] != spart[b]) return spart[a] < spart[b];
        const Key ka = skey[a], kb = skey[b];
        for (uint32_t r = 0;; ++r) {
            uint64_t ca = sort_chunk(ka, bytes, r), cb = sort_chunk(kb, bytes, r);
            if (ca != cb) return ca < cb;
            if ((ca & 0xFF) < 8) break;
        }
        return sval[a] < sval[b];
    };
    for (uint32_t k = 2; k <= np2; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < np2; i += blockDim.x) {
                uint32_t ixj = i ^ j;
                if (ixj > i) {
                    bool up = (i & k) == 0;
                    uint32_t pa = posv[i], pb = posv[ixj];
                    if ((pa > pb) == up) { posv[i] = pb; posv[ixj] = pa; }
                    uint32_t a = order[i], b = order[ixj];
                    bool swap = up ? less(b, a) : less(a, b);
                    if (swap) { order[i] = (uint16_t)b; order[ixj] = (uint16_t)a; }
                }
            }
            __syncthreads();
        }
    const uint32_t fin = ctl->npasses & 1;
    for (uint32_t i = threadIdx.x; i < nt; i += blockDim.x) bufs.val[fin][posv[i]] = sval[order[i]];
}

// General path (host-driven): bitonic steps over the tie list in global memory.
__global__ void __launch_bounds__(256) tie_pad_kernel(TieItem *ties, uint32_t *posv, uint32_t nt,
                                                      uint32_t np2)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np2) return;
    if (i >= nt) { ties[i].pos = 0xFFFFFFFFu; ties[i].val = 0; }
    posv[i] = ties[i].pos;
}

__global__ void __launch_bounds__(256) tie_bitonic_step_kernel(TieItem *items, uint32_t *posv,
                                                               uint32_t np2, uint32_t k, uint32_t j,
                                                               EntryArrays ent, const uint8_t *bytes)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np2) return;
    uint32_t ixj = i ^ j;
    if (ixj <= i) return;
    bool up = (i & k) == 0;
    uint32_t pa = posv[i], pb = posv[ixj];
    if ((pa > pb) == up) { posv[i] = pb; posv[ixj] = pa; }
    TieItem a = items[i], b = items[ixj];
    bool swap = up ? tie_less(b, a, ent, bytes) : tie_less(a, b, ent, bytes);
    if (swap) { items[i] = b; items[ixj] = a; }
}

__global__ void __launch_bounds__(256) tie_writeback_kernel(SortBufs bufs, const TieItem *items,
                                                            const uint32_t *posv, uint32_t nt, Ctl *ctl)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nt) bufs.val[ctl->npasses & 1][posv[i]] = items[i].val;
}

// ------------------------------------------------------------------ scans (3 kernels)
// exclusive scan of a uint32 array of ctl-> length into uint64; `which` selects the
// length and where the total goes: 0 -> n entries, total to n_out; 1 -> n_out rows,
// total to key_bytes.
__device__ __forceinline__ uint64_t scan_len(const Ctl *ctl, int which) { return which == 0 ? ctl->n : ctl->n_out; }

__global__ void __launch_bounds__(256) scan_partial_kernel(const uint32_t *in, unsigned long long *partial,
                                                           const Ctl *ctl, int which)
{
    const uint64_t n = scan_len(ctl, which);
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    if (base >= n) return;
    __shared__ unsigned long long wsum[8];
    unsigned long long s = 0;
    for (uint32_t i = threadIdx.x; i < SCAN_TILE; i += 256) { uint64_t idx = base + i; if (idx < n) s += in[idx]; }
    for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long t = 0; for (int w = 0; w < 8; ++w) t += wsum[w]; partial[blockIdx.x] = t; }
}

__global__ void __launch_bounds__(1024) scan_spine_kernel(unsigned long long *partial, Ctl *ctl, int which)
{
    const uint64_t n = scan_len(ctl, which);
    const uint32_t nt = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < nt; base += 1024) {
        uint32_t i = base + threadIdx.x;
        unsigned long long v = i < nt ? partial[i] : 0, incl = v;
        for (int d = 1; d < 32; d <<= 1) { unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        unsigned long long woff = 0;
        for (uint32_t w = 0; w < warp; ++w) woff += wsum[w];
        unsigned long long c = carry;
        if (i < nt) partial[i] = c + woff + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + woff + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) { if (which == 0) ctl->n_out = carry; else ctl->key_bytes = carry; }
}

__global__ void __launch_bounds__(256) scan_final_kernel(const uint32_t *in, const unsigned long long *partial,
                                                         unsigned long long *out, const Ctl *ctl, int which)
{
    const uint64_t n = scan_len(ctl, which);
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    if (base >= n) { return; }
    __shared__ unsigned long long wsum[8];
    // thread t handles 16 consecutive items
    const uint64_t first = base + (uint64_t)threadIdx.x * 16;
    uint32_t v[16];
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) { uint64_t idx = first + i; v[i] = idx < n ? in[idx] : 0; s += v[i]; }
    unsigned long long incl = s;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = 1; d < 32; d <<= 1) { unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    unsigned long long off = partial[blockIdx.x] + incl - s;
    for (uint32_t w = 0; w < warp; ++w) off += wsum[w];
#pragma unroll
    for (int i = 0; i < 16; ++i) { uint64_t idx = first + i; if (idx < n) out[idx] = off; off += v[i]; }
}

// ------------------------------------------------------------------ K6 / K7
// head[i] = 1 when sorted position i starts a new key
__global__ void __launch_bounds__(256) head_mark_kernel(SortBufs bufs, EntryArrays ent, const uint8_t *bytes,
                                                        uint32_t *head, const Ctl *ctl)
{
    const uint64_t n = ctl->n;
    const uint32_t fin = ctl->npasses & 1;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t h = 1;
        if (i > 0 && bufs.hi[fin][i] == bufs.hi[fin][i - 1] && bufs.lo[fin][i] == bufs.lo[fin][i - 1])
            h = same_key(ent.key(bufs.val[fin][i]), ent.key(bufs.val[fin][i - 1]), bytes) ? 0u : 1u;
        head[i] = h;
    }
}

struct RowArrays {
    uint32_t *part;
    unsigned long long *count;
    unsigned long long *key_off; // [n_out + 1]
    uint8_t *key_bytes;
    unsigned long long *k0, *k1; // key words of the row (internal)
    uint32_t *len;
    unsigned long long *part_begin; // [R + 1]
};

// every head sums its run of equal keys and writes the row
__global__ void __launch_bounds__(256) reduce_rows_kernel(SortBufs bufs, EntryArrays ent, const uint32_t *head,
                                                          const unsigned long long *out_idx, RowArrays rows,
                                                          const Ctl *ctl)
{
    const uint64_t n = ctl->n;
    const uint32_t fin = ctl->npasses & 1;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (!head[i]) continue;
        uint32_t v = bufs.val[fin][i];
        unsigned long long c = ent.cnt[v];
        for (uint64_t j = i + 1; j < n && !head[j]; ++j) c += ent.cnt[bufs.val[fin][j]];
        uint64_t o = out_idx[i];
        const Key key = ent.key(v);
        rows.part[o] = ent.part[v];
        rows.count[o] = c;
        rows.k0[o] = key.k0; rows.k1[o] = key.k1;
        rows.len[o] = key_len(key);
    }
}

__global__ void __launch_bounds__(256) write_key_bytes_kernel(RowArrays rows, const uint8_t *bytes,
                                                              uint32_t num_reducers, const Ctl *ctl)
{
    if (ctl->err & ERR_KEYS_FULL) return; // rows or key bytes do not fit: reported by the host
    const uint64_t n = ctl->n_out;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t o = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; o <= n; o += stride) {
        // partition starts: rows are partition-sorted
        uint32_t p_prev = o == 0 ? 0u : rows.part[o - 1] + 1u;
        uint32_t p_cur = o == n ? num_reducers : rows.part[o];
        for (uint32_t p = p_prev; p <= p_cur && p <= num_reducers; ++p) rows.part_begin[p] = o;
        if (o == n) { rows.key_off[n] = ctl->key_bytes; break; }
        Key key; key.k0 = rows.k0[o]; key.k1 = rows.k1[o];
        uint8_t *dst = rows.key_bytes + rows.key_off[o];
        uint32_t len = rows.len[o];
        if (!key_is_long(key)) {
            for (uint32_t i = 0; i < len && i < 8; ++i) dst[i] = (uint8_t)(key.k0 >> (8 * i));
            for (uint32_t i = 8; i < len; ++i) dst[i] = (uint8_t)(key.k1 >> (8 * (i - 8)));
        } else {
            uint64_t off = key_off(key);
            for (uint32_t i = 0; i < len; i += 8) {
                uint64_t w = load8(bytes, off + i);
                uint32_t m = len - i < 8 ? len - i : 8;
                for (uint32_t j = 0; j < m; ++j) dst[i + j] = (uint8_t)(w >> (8 * j));
            }
        }
    }
}

} // namespace mr
