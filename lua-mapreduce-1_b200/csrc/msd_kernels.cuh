// msd_kernels.cuh -- the reduce side for jobs of up to a few million distinct keys, built for latency:
// three launches after the table read-out instead of eleven.
//
// Replaces utils.keys_sorted (utils.lua:123-128), utils.merge_iterator + heap.lua (utils.lua:206-271),
// merge_min_keys (utils.lua:232-247), the reduce loop job.lua:264-284 and the result write
// job.lua:272-287 -- the same reference code the LSD path of reduce_kernels.cuh replaces; that path
// stays for what this one declines (see "fallback").
//
//   msd_hist_kernel     the sort word (partition, 7 key bytes, min(len, 8)) is read as 9 fields; every
//                       key byte is replaced by its rank among the byte values that occur at its
//                       position (dense codes whatever the alphabet), and the bin of an item is the
//                       mixed-radix number of its leading fields, refined until a bin holds ~8 items
//                       (at most 2^15 bins): one histogram over all items.
//   msd_scatter_kernel  bin starts (every CTA scans the histogram for itself), then every item goes to
//                       its bin as {packed low bits, entry index, bin}: items of one bin lie together,
//                       bins in key order.
//   msd_unit_kernel     units of consecutive whole bins (about 512 items, never more than 1024) are
//                       sorted in SHARED memory (stable LSD passes on the packed low bits and the bin
//                       number), keys that agree in their first 7 bytes are ordered by their later
//                       bytes, equal keys (the same key from several ranks) are merged, and the rows --
//                       partition, count, key offsets, packed key bytes, partition starts -- are
//                       written straight from the unit, numbered by a decoupled look-back over units.
//
// Fallback: the path declines (Ctl::msd_fallback, nothing written that the LSD path would not
// overwrite) when the packed low bits exceed 64, a bin holds more than UNIT_CAP - UNIT_W items
// (many more keys than 2^15 bins can spread, or keys that only differ late), or a group of keys
// equal in their first 7 bytes has more than MSD_TIE_GROUP members (copies of a key from several ranks count).  mr_job_sync then runs the LSD
// path, and so do the context's later jobs (every job is then complete without a host round trip) until
// its owner ranks are set anew.
#pragma once
#include "reduce_kernels.cuh"

namespace mr {

constexpr int MSD_MAX_BITS = 15;
constexpr uint32_t MSD_MAX_BINS = 1u << MSD_MAX_BITS;
#define MSD_PAD(b) ((b) + ((b) >> 5))                              // index of bin b in the scatter kernel's shared-memory table
constexpr uint32_t MSD_SCATTER_SMEM = (MSD_MAX_BINS + MSD_MAX_BINS / 32 + 8) * 4;
constexpr uint32_t MSD_TARGET_BIN = 8;       // items per bin aimed at (a bin may hold up to the unit size - UNIT_W: room for 64x skew)
constexpr uint32_t UNIT_W = 512;             // a unit owns the bins that start inside its 512 positions
// ... so it holds fewer than UNIT_W + largest bin items.  Two sizes of unit: jobs without a shuffle
// (every key once) get four 1024-item units per SM; after a shuffle a key arrives from every rank, bins
// are that much fuller where the vocabulary is crowded, and the units hold 1536 (three per SM).
constexpr uint32_t UNIT_CAP_DISTINCT = 1024, UNIT_CAP_SHUFFLED = 1536;
constexpr int UNIT_THREADS = 256;
constexpr int UNIT_WARPS = UNIT_THREADS / 32;
#ifndef MRGPU_COUNT_SORT_MAX
#define MRGPU_COUNT_SORT_MAX 48
#endif
constexpr uint32_t UNIT_COUNT_SORT_MAX = MRGPU_COUNT_SORT_MAX;  // units whose largest bin is at most this: bins ordered by counting inside the bin (b compares per item); fuller ones: radix passes (~14 compares' worth per pass)
constexpr uint32_t MSD_TIE_GROUP = 64;        // keys equal in their first 7 bytes (copies of one key included) a unit orders itself

enum : uint32_t { MSD_FB_WIDE = 1u, MSD_FB_BIN = 4u, MSD_FB_TIES = 8u };

// histogram, scatter cursors and bin starts; two of these alternate by job so that each job finds
// its own zeroed (the histogram kernel of a job clears the other one)
struct MsdBins {
    unsigned int hist[MSD_MAX_BINS];
    unsigned int bin_start[MSD_MAX_BINS + 1];
    unsigned int unit_bin[MSD_MAX_BINS + 2];   // first bin that starts at or after unit * UNIT_W (n <= bins * UNIT_W when no bin is too large, so units <= bins)
    unsigned int unit_ticket;
};
static_assert(sizeof(MsdBins) % 16 == 0, "MsdBins is cleared 16 bytes at a time");

// what the scatter leaves in a bin
struct __align__(16) MsdItem {
    unsigned long long low;  // packed varying bits below the digit
    uint32_t val;            // entry index
    uint32_t bin;            // bin | key length << 15 (jobs of distinct keys) | MSD_FLAG8 when the key has 8 or more bytes
};
constexpr uint32_t MSD_FLAG8 = 1u << 31;
constexpr uint32_t MSD_BIN_MASK = MSD_MAX_BINS - 1;      // bits 0..14: the bin
constexpr uint32_t MSD_LEN_SHIFT = 15, MSD_LEN_MAX = 0xFFFFu; // bits 15..30: min(key length, 65535), only in jobs of distinct keys
static_assert(sizeof(MsdItem) == sizeof(SortItem), "the scattered items reuse the second sort buffer");

__host__ __device__ __forceinline__ uint32_t popc32_hd(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
__host__ __device__ __forceinline__ uint32_t bits_for(uint32_t card) // bits that hold the ranks 0..card-1
{
    return card <= 1 ? 0u : 32u - (uint32_t)clz32(card - 1);
}

// The plan from the occupancy masks (occ[p][8] = 256-bit mask of the byte values at key byte p), the
// partition count and the number of items.
__host__ __device__ inline void msd_build_plan(const unsigned int (*occ)[8], uint32_t part_min, uint32_t part_max,
                                               unsigned long long n, MsdPlan *p)
{
    // the partitions PRESENT, not num_reducers: a rank of many owns a key range of one or two partitions,
    // and bins spent on the others would leave its own too crowded
    p->part_base = part_min <= part_max ? part_min : 0;
    p->card[0] = part_min <= part_max ? part_max - part_min + 1 : 1;
    for (int f = 0; f < 7; ++f) {
        uint32_t c = 0;
        for (int w = 0; w < 8; ++w) c += popc32_hd(occ[f][w]);
        p->card[1 + f] = c ? c : 1;
    }
    p->card[8] = 9;
    uint32_t low_all = 0;
    for (int f = 0; f < MSD_FIELDS; ++f) { p->width[f] = (uint8_t)bits_for(p->card[f]); low_all += p->width[f]; }
    unsigned long long T = n / MSD_TARGET_BIN;
    if (T < 1) T = 1;
    if (T > MSD_MAX_BINS) T = MSD_MAX_BINS;
    unsigned long long P = 1;
    uint32_t k = 0;
    while (k < (uint32_t)MSD_FIELDS && P * p->card[k] <= T) { P *= p->card[k]; ++k; }
    unsigned long long e = 1;
    if (k < (uint32_t)MSD_FIELDS) {
        e = T / P;
        if (e < 1) e = 1;
        if (e > p->card[k]) e = p->card[k];
    }
    p->k = k;
    p->e = (uint32_t)e;
    p->nbins = (uint32_t)(P * e);
    uint32_t low_bits = low_all;
    for (uint32_t f = 0; f < k; ++f) low_bits -= p->width[f];
    p->low_bits = low_bits;
    p->fallback = low_bits > 64 ? MSD_FB_WIDE : 0;
    p->pad[0] = p->pad[1] = p->pad[2] = 0;
    p->pad2 = 0;
}

// rank of byte value v under a 256-bit occupancy mask: how many occurring values are smaller
__host__ __device__ __forceinline__ uint32_t msd_rank_of(const unsigned int *occ8, uint32_t v)
{
    uint32_t r = 0;
    for (uint32_t w = 0; w < (v >> 5); ++w) r += popc32_hd(occ8[w]);
    return r + popc32_hd(occ8[v >> 5] & ((1u << (v & 31)) - 1u));
}

// bin and packed low word of one sort word (hi = partition, lo = 7 key bytes big-endian << 8 | min(len, 8))
// lutk (optional): for a plan whose split field k is a key byte, lutk[v] = rank(v) * e / card[k] -- the
// same value the division gives, looked up
__host__ __device__ __forceinline__ void msd_extract(const MsdPlan &p, const unsigned char (*lut)[256], uint32_t hi,
                                                     unsigned long long lo, uint32_t &bin, unsigned long long &low,
                                                     const unsigned char *lutk = nullptr)
{
    uint32_t rank[MSD_FIELDS];
    rank[0] = hi - p.part_base;
#pragma unroll
    for (int f = 0; f < 7; ++f) rank[1 + f] = lut[f][(uint32_t)(lo >> (56 - 8 * f)) & 0xFFu];
    rank[8] = (uint32_t)lo & 0xFFu;
    uint32_t b = 0;
    unsigned long long l = 0;
#pragma unroll
    for (int f = 0; f < MSD_FIELDS; ++f) {
        if ((uint32_t)f < p.k) b = b * p.card[f] + rank[f];
        else {
            if ((uint32_t)f == p.k) b = b * p.e + (lutk && f >= 1 && f <= 7 ? (uint32_t)lutk[(uint32_t)(lo >> (56 - 8 * (f - 1))) & 0xFFu]
                                                                           : (uint32_t)(((unsigned long long)rank[f] * p.e) / p.card[f]));
            l = (l << p.width[f]) | rank[f];
        }
    }
    bin = b;
    low = l;
}

// 8-bit digit `pass` of the virtual number (local bin : low), least significant first
__host__ __device__ __forceinline__ uint32_t msd_pass_digit(unsigned long long low, uint32_t lbin, uint32_t low_bits, uint32_t pass)
{
    const uint32_t b0 = pass * 8;
    unsigned long long v;
    if (b0 >= low_bits) v = (unsigned long long)lbin >> (b0 - low_bits);
    else {
        v = low >> b0;                                   // b0 < low_bits <= 64
        if (b0 + 8 > low_bits) v |= (unsigned long long)lbin << (low_bits - b0);
    }
    return (uint32_t)v & 0xFFu;
}

// Two inline keys (<= 15 bytes) that agree in bytes 0..6 and have 8 or more: their order is that of
// (bytes 7..14 as one big-endian word, zero-padded; then the length) -- what key_less / sort_chunk give
// round by round, in one compare (checked on the host: tests/test_device_helpers_on_host.py).
__host__ __device__ __forceinline__ unsigned long long inline_tail_word(const Key &k)
{
    return bswap64((k.k0 >> 56) | ((k.k1 & LOW56) << 8));
}
__host__ __device__ __forceinline__ bool inline_tail_less(const Key &a, const Key &b, bool a_first_if_equal)
{
    const unsigned long long ra = inline_tail_word(a), rb = inline_tail_word(b);
    const uint32_t la = (uint32_t)(a.k1 >> 56), lb = (uint32_t)(b.k1 >> 56);
    return ra < rb || (ra == rb && (la < lb || (la == lb && a_first_if_equal)));
}

// ------------------------------------------------------------------ histogram
// tag of an item, left by the histogram kernel for the scatter: bin | 8+ bytes flag << 15 | place inside the bin << 16
constexpr uint32_t MSD_TAG_FLAG8 = 1u << 15;

__global__ void __launch_bounds__(256) msd_hist_kernel(const SortItem *items, unsigned long long max_keys, uint32_t num_reducers,
                                                       Ctl *ctl, MsdBins *mb, MsdBins *mb_next, uint32_t *tags, RowsStatus st,
                                                       uint32_t st_words)
{
    __shared__ MsdPlan plan;
    __shared__ unsigned int s_occ[7][8];
    __shared__ unsigned char lut[7][256];
    __shared__ unsigned char lutk[256];
    unsigned long long n = ctl->n;
    const bool over = n > max_keys;
    if (over) n = max_keys;
    // the other job's bins, and the look-back words of the units to come
    {
        const uint64_t g0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, gs = (uint64_t)gridDim.x * blockDim.x;
        uint4 *z = (uint4 *)mb_next;
        for (uint64_t i = g0; i < sizeof(MsdBins) / 16; i += gs) z[i] = make_uint4(0, 0, 0, 0);
        const uint64_t units = (n + UNIT_W - 1) / UNIT_W + 1;
        for (uint64_t i = g0; i < units && i < st_words; i += gs) { st.heads[i] = 0; st.bytes_agg[i] = 0; st.bytes_inc[i] = 0; }
    }
    if (threadIdx.x < 56) (&s_occ[0][0])[threadIdx.x] = (&ctl->occ[0][0])[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) msd_build_plan(s_occ, ctl->part_min, ctl->part_max, n, &plan);
    for (int f = 0; f < 7; ++f) lut[f][threadIdx.x] = (unsigned char)msd_rank_of(s_occ[f], threadIdx.x);   // 256 threads
    __syncthreads();
    {
        const uint32_t k = plan.k;
        lutk[threadIdx.x] = (k >= 1 && k <= 7) ? (unsigned char)(((uint32_t)lut[k - 1][threadIdx.x] * plan.e) / plan.card[k]) : 0;
    }
    __syncthreads();
    if (blockIdx.x == 0) {
        ctl->msd_lutk[threadIdx.x] = lutk[threadIdx.x];
        uint32_t *dst = (uint32_t *)&ctl->msd_plan;
        const uint32_t *src = (const uint32_t *)&plan;
        for (uint32_t i = threadIdx.x; i < sizeof(MsdPlan) / 4; i += blockDim.x) dst[i] = src[i];
        for (int f = 0; f < 7; ++f) ctl->msd_lut[f][threadIdx.x] = lut[f][threadIdx.x];
        if (threadIdx.x == 0) {
            ctl->msd_fallback = plan.fallback;
            if (over) { ctl->n = max_keys; atomicOr(&ctl->err, ERR_KEYS_FULL); }
        }
    }
    if (plan.fallback || n == 0) return;
    // one atomic per item: counts the bin and tells the item its place inside it (any order will do:
    // the unit sorts the bin).  Four items per thread in flight.
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
        SortItem it[4];
        uint32_t d[4], r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint64_t i = i0 + q * stride;
            if (i < n) it[q] = items[i];
            else { it[q].lo = 0; it[q].hi = 0; it[q].val = 0; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned long long low;
            msd_extract(plan, lut, it[q].hi, it[q].lo, d[q], low, lutk);
            r[q] = i0 + q * stride < n ? atomicAdd(&mb->hist[d[q]], 1u) : 0u;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint64_t i = i0 + q * stride;
            if (i < n) tags[i] = d[q] | ((it[q].lo & 0xFFu) == 8u ? MSD_TAG_FLAG8 : 0u) | min(r[q], 0xFFFFu) << 16;
        }
    }
}

// ------------------------------------------------------------------ scatter
__global__ void __launch_bounds__(1024) msd_scatter_kernel(const SortItem *items, const uint32_t *tags, MsdItem *out, Ctl *ctl,
                                                           MsdBins *mb, const uint32_t *lens, uint32_t max_bin)
{
    // [bins + 1], one pad word per 32 (MSD_PAD): a thread's run of consecutive bins would otherwise hit one
    // bank with every lane (with 30 K bins -- an owner's share after a shuffle -- the scan took 25 us)
    extern __shared__ unsigned int msd_start[];
    __shared__ unsigned int wsum[32];
    __shared__ unsigned int s_big;
    __shared__ MsdPlan plan;
    __shared__ unsigned char lut[7][256];
    if (ctl->msd_fallback) return;
    const unsigned long long n = ctl->n;
    if (n == 0) return;
    for (uint32_t i = threadIdx.x; i < sizeof(MsdPlan) / 4; i += blockDim.x) ((uint32_t *)&plan)[i] = ((const uint32_t *)&ctl->msd_plan)[i];
    for (uint32_t i = threadIdx.x; i < 7 * 256 / 4; i += blockDim.x) ((uint32_t *)&lut[0][0])[i] = ((const uint32_t *)&ctl->msd_lut[0][0])[i];
    if (threadIdx.x == 0) s_big = 0;
    __syncthreads();
    const uint32_t bins = plan.nbins;
    // exclusive scan of the histogram: 1024 threads x (bins / 1024) consecutive bins each
    const uint32_t per = (bins + 1023) / 1024;
    const uint32_t b0 = threadIdx.x * per;
    // the histogram comes in with coalesced loads; each thread then takes its consecutive bins from shared memory
    for (uint32_t b = threadIdx.x; b < bins; b += 1024) msd_start[MSD_PAD(b)] = mb->hist[b];
    __syncthreads();
    unsigned int c[MSD_MAX_BINS / 1024];
    unsigned int sum = 0, big = 0;
#pragma unroll
    for (uint32_t k = 0; k < MSD_MAX_BINS / 1024; ++k) {
        c[k] = (k < per && b0 + k < bins) ? msd_start[MSD_PAD(b0 + k)] : 0;
        sum += c[k];
        big = max(big, c[k]);
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
    if (lane == 31) wsum[warp] = incl;
    if (big > max_bin) atomicOr(&s_big, 1u);
    __syncthreads();
    unsigned int woff = 0;
    for (uint32_t w = 0; w < warp; ++w) woff += wsum[w];
    unsigned int run = woff + incl - sum;
#pragma unroll
    for (uint32_t k = 0; k < MSD_MAX_BINS / 1024; ++k)
        if (k < per && b0 + k < bins) { msd_start[MSD_PAD(b0 + k)] = run; run += c[k]; }
    if (threadIdx.x == 0) msd_start[MSD_PAD(bins)] = (unsigned int)n;
    __syncthreads();
    if (s_big) {                                    // every CTA sees the same histogram and takes the same way out
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(&ctl->msd_fallback, MSD_FB_BIN);
        return;
    }
    {
        const uint32_t g0 = blockIdx.x * blockDim.x + threadIdx.x, gs = gridDim.x * blockDim.x;
        for (uint32_t i = g0; i <= bins; i += gs) mb->bin_start[i] = msd_start[MSD_PAD(i)];
        // unit u owns the bins from the first one that starts at or after u * UNIT_W: bin b is that
        // first bin for every u with start[b-1] < u * UNIT_W <= start[b]
        const uint32_t nunits = (uint32_t)((n + UNIT_W - 1) / UNIT_W);
        for (uint32_t b = g0; b <= bins; b += gs) {
            const uint32_t hi = msd_start[MSD_PAD(b)] / UNIT_W;                                  // last u with u * W <= start[b]
            const uint32_t lo = b == 0 ? 0u : (msd_start[MSD_PAD(b - 1)] + UNIT_W) / UNIT_W;      // first u with u * W > start[b-1] ... (start[b-1] / W) + 1
            for (uint32_t u = lo; u <= hi && u <= nunits; ++u) mb->unit_bin[u] = b;
        }
    }
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
        SortItem it[4];
        uint32_t tg[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint64_t i = i0 + q * stride;
            if (i < n) { it[q] = items[i]; tg[q] = tags[i]; }
            else { it[q].lo = 0; it[q].hi = 0; it[q].val = 0; tg[q] = 0; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (i0 + q * stride >= n) continue;
            uint32_t d;
            MsdItem o;
            msd_extract(plan, lut, it[q].hi, it[q].lo, d, o.low, nullptr);   // the low word only; the bin travels in the tag
            d = tg[q] & 0x7FFFu;
            o.val = it[q].val;
            o.bin = d | ((tg[q] & MSD_TAG_FLAG8) ? MSD_FLAG8 : 0u);
            if (lens) o.bin |= min(lens[i0 + q * stride], MSD_LEN_MAX) << MSD_LEN_SHIFT;   // jobs of distinct keys: the key's length rides along
            out[msd_start[MSD_PAD(d)] + (tg[q] >> 16)] = o;
        }
    }
}

// ------------------------------------------------------------------ units: sort, merge, rows
template <uint32_t UNIT_CAP>
struct __align__(16) UnitSmem {
    MsdItem buf[2][UNIT_CAP];                    // ping-pong of the sort; afterwards buf[other] holds the keys
    unsigned int whist[UNIT_CAP * 2 > UNIT_WARPS * RADIX_BINS ? UNIT_CAP * 2 : UNIT_WARPS * RADIX_BINS]; // bin extents, per-warp digit counts [UNIT_WARPS][RADIX_BINS], afterwards the values of the items
    unsigned int bin_excl[RADIX_BINS];
    unsigned int scan_tmp[RADIX_BINS / 32];
    unsigned int wh[UNIT_WARPS];
    unsigned long long wb[UNIT_WARPS];
    unsigned long long prefix_h, prefix_b;
    unsigned int prefix_part, unit, first, count, dfirst, fallback, maxbin, inexact;
    unsigned long long sumlen;
    unsigned char iflag[UNIT_CAP];               // per item: continues its predecessor's word / another key; later: starts a row
};

#ifdef MRGPU_UNIT_TIMING   // build with -DMRGPU_UNIT_TIMING=<unit>: globaltimer stamps of that unit's phases in Ctl::dbg
#define UT(k) do { if (threadIdx.x == 0 && sm.unit == (MRGPU_UNIT_TIMING)) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); ctl->dbg[k] = (long long)t_; } } while (0)
#define UT0 do { if (threadIdx.x == 0 && sm.unit == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); ctl->dbg[15] = (long long)t_; } } while (0)
#else
#define UT(k) do { } while (0)
#define UT0 do { } while (0)
#endif

template <uint32_t UNIT_CAP, int MIN_CTAS>
__global__ void __launch_bounds__(UNIT_THREADS, MIN_CTAS) msd_unit_kernel(const MsdItem *in, EntryArrays ent, const uint8_t *bytes,
                                                                RowArrays rows, uint32_t num_reducers, uint32_t op,
                                                                unsigned long long key_bytes_cap, RowsStatus st, Ctl *ctl,
                                                                MsdBins *mb, bool distinct)
{
    extern __shared__ __align__(16) uint8_t unit_smem_raw[];
    constexpr int UNIT_ITEMS = UNIT_CAP / UNIT_THREADS;
    using Smem = UnitSmem<UNIT_CAP>;
    Smem &sm = *(Smem *)unit_smem_raw;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, lt = lanemask_lt();
    const unsigned long long n = ctl->n;
    if (tid == 0) {
        sm.unit = atomicAdd(&mb->unit_ticket, 1u);   // units are taken in order: a unit only waits for started ones
        sm.fallback = ctl->msd_fallback;
    }
    __syncthreads();
    const uint32_t unit = sm.unit;
    UT0; UT(0);
    if (sm.fallback) return;
    if (n == 0) {
        if (unit == 0) {
            for (uint32_t p = tid; p <= num_reducers; p += UNIT_THREADS) rows.part_begin[p] = 0;
            if (tid == 0) { ctl->n_out = 0; ctl->key_bytes = 0; rows.key_off[0] = 0; }
        }
        return;
    }
    const uint32_t nunits = (uint32_t)((n + UNIT_W - 1) / UNIT_W);
    if (unit >= nunits) return;
    const uint32_t bins = ctl->msd_plan.nbins, low_bits = ctl->msd_plan.low_bits;
    if (tid == 0) {
        // bins whose start lies in [unit * W, (unit + 1) * W) (table left by the scatter kernel)
        const uint32_t ba = mb->unit_bin[unit];
        const uint32_t be = unit + 1 >= nunits ? bins : mb->unit_bin[unit + 1];
        const uint32_t s0 = mb->bin_start[ba];
        sm.first = s0;
        sm.count = mb->bin_start[be] - s0;
        sm.dfirst = ba;
        sm.maxbin = 0;
        sm.inexact = 0;
        sm.sumlen = 0;
    }
    __syncthreads();
    const uint32_t first = sm.first, m = sm.count, dfirst = sm.dfirst;   // m < UNIT_CAP (checked by the scatter kernel)
    UT(1);

    // ---- load; every item learns where its bin lies inside the unit
    unsigned int *ext = &sm.whist[0];            // [i] = bin begin | bin size << 16 (free until the passes / the counts)
    {
        uint32_t big = 0;
        unsigned long long lsum = 0;
        bool inexact = false;
        for (uint32_t i = tid; i < m; i += UNIT_THREADS) {
            MsdItem it = in[first + i];
            const uint32_t d = it.bin & MSD_BIN_MASK;
            const uint32_t l16 = (it.bin >> MSD_LEN_SHIFT) & MSD_LEN_MAX;
            lsum += l16;
            inexact |= l16 == MSD_LEN_MAX || l16 == 0;
            const uint32_t b0 = mb->bin_start[d] - first, sz = mb->bin_start[d + 1] - mb->bin_start[d];
            it.bin = (it.bin & MSD_FLAG8) | (d - dfirst);
            sm.buf[0][i] = it;
            ext[i] = b0 | sz << 16;
            big = max(big, sz);
        }
        for (int d = 16; d; d >>= 1) {
            big = max(big, __shfl_xor_sync(0xFFFFFFFFu, big, d));
            lsum += __shfl_xor_sync(0xFFFFFFFFu, lsum, d);
        }
        if (lane == 0 && big) atomicMax(&sm.maxbin, big);
        if (lane == 0 && lsum) atomicAdd(&sm.sumlen, lsum);
        if (inexact) sm.inexact = 1;
    }
    uint32_t maxl = 0;
    if (m) maxl = (in[first + m - 1].bin & MSD_BIN_MASK) - dfirst;           // bins ascend with the position
    uint32_t lbits = 0;
    while (lbits < 32 && (1u << lbits) <= maxl) ++lbits;
    const uint32_t npass = (low_bits + lbits + 7) / 8;
    __syncthreads();
    UT(2);

    // ---- look-back over units: rows before, key bytes before, partition of the last row before.
    // Two self-validating words per unit, each led by the same 2-bit state (1 aggregate, 2 inclusive):
    //   A = state << 62 | (last partition + 1) << 32 | rows      (0 in the partition field = no row)
    //   B = state << 62 | key bytes
    // A reader takes a pair only when both states agree, so no fence orders the two stores; 128
    // predecessors are read per round trip.
    constexpr unsigned long long F_AGG = 1ull << 62, F_INC = 2ull << 62, F_VAL = (1ull << 62) - 1;
    auto lookback = [&](unsigned long long th_, unsigned long long tb_, unsigned long long last_) {   // warp 0
        {
            volatile unsigned long long *vA = st.heads, *vB = st.bytes_agg;
            unsigned long long ph = 0, pb = 0, pl = 0;
            if (unit == 0) {
                if (lane == 0) { vB[0] = F_INC | tb_; vA[0] = F_INC | last_ << 32 | th_; }
            } else {
                if (lane == 0) { vB[unit] = F_AGG | tb_; vA[unit] = F_AGG | last_ << 32 | th_; }
                int t = (int)unit - 1;
                bool done = false;
                for (uint32_t polls = 0; !done; ++polls) {
                    if (polls > LOOKBACK_SPIN_LIMIT) { if (lane == 0) atomicOr(&ctl->err, ERR_LOOKBACK_TIMEOUT); break; }
                    unsigned long long wa[4], wb[4];
    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int idx = t - q * 32 - (int)lane;
                        wa[q] = idx >= 0 ? vA[idx] : F_INC;
                        wb[q] = idx >= 0 ? vB[idx] : F_INC;
                    }
    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t fa = (uint32_t)(wa[q] >> 62), fb = (uint32_t)(wb[q] >> 62);
                        const bool ok = fa == fb && fa != 0;
                        const uint32_t notready = __ballot_sync(0xFFFFFFFFu, !ok);
                        const uint32_t incm = __ballot_sync(0xFFFFFFFFu, ok && fa == 2);
                        const uint32_t stop = incm ? (uint32_t)__ffs((int)incm) - 1u : 31u;
                        const uint32_t use = stop == 31 ? 0xFFFFFFFFu : ((2u << stop) - 1u);
                        if (notready & use) break;                 // a needed predecessor has not published: read again from here
                        unsigned long long bb = 0, h = 0, l = 0;
                        if ((use >> lane) & 1u) {
                            h = wa[q] & 0xFFFFFFFFull;
                            l = (wa[q] >> 32) & 0x3FFFFFFFull;
                            bb = wb[q] & F_VAL;
                        }
                        for (int d = 16; d; d >>= 1) {
                            h += __shfl_xor_sync(0xFFFFFFFFu, h, d);
                            bb += __shfl_xor_sync(0xFFFFFFFFu, bb, d);
                            l = max(l, __shfl_xor_sync(0xFFFFFFFFu, l, d));   // partitions ascend with the units
                        }
                        ph += h; pb += bb; pl = max(pl, l);
                        if (incm) { done = true; break; }
                        t -= 32;
                    }
                }
                if (lane == 0) { vB[unit] = F_INC | (pb + tb_); vA[unit] = F_INC | max(pl, last_) << 32 | (ph + th_); }
            }
            if (lane == 0) { sm.prefix_h = ph; sm.prefix_b = pb; sm.prefix_part = (unsigned int)pl; }
        }
    };
    // Jobs without a shuffle hold every key once: each item will be a row, its key length came along
    // with it, and the partition is the leading field of the bin -- so the unit publishes its rows and
    // bytes NOW and the wait for the units before it overlaps the sorting below.
    const uint32_t plan_k = ctl->msd_plan.k;
    const bool early = distinct && plan_k >= 1 && !sm.inexact && !sm.fallback;
    if (early && warp == 0) {
        const uint32_t bpp = bins / ctl->msd_plan.card[0];
        lookback(m, sm.sumlen, m ? (unsigned long long)((dfirst + maxl) / bpp + ctl->msd_plan.part_base + 1u) : 0ull);
    }

    uint32_t cur = 0;
    const bool small_bins = sm.maxbin <= UNIT_COUNT_SORT_MAX;
    if (small_bins) {
        // ---- the usual case, bins of a few dozen items: every item counts the items of its own bin
        // that come before it (smaller low word, or equal and earlier) -- one pass, one barrier
        const bool narrow = low_bits <= 32;            // (the usual case: the bin already fixes the leading fields)
        for (uint32_t i = tid; i < m; i += UNIT_THREADS) {
            const MsdItem it = sm.buf[0][i];
            const uint32_t b0 = ext[i] & 0xFFFFu, sz = ext[i] >> 16;
            uint32_t rank = 0;
            if (narrow) {
                const uint32_t mine = (uint32_t)it.low;
                const uint32_t *lo32 = (const uint32_t *)&sm.buf[0][0];   // low word of item j at [4 * j]
                uint32_t j = b0;
                for (; j + 4 <= b0 + sz; j += 4) {
                    const uint32_t l0 = lo32[4 * j], l1 = lo32[4 * j + 4], l2 = lo32[4 * j + 8], l3 = lo32[4 * j + 12];
                    rank += (l0 < mine || (l0 == mine && j < i)) + (l1 < mine || (l1 == mine && j + 1 < i)) +
                            (l2 < mine || (l2 == mine && j + 2 < i)) + (l3 < mine || (l3 == mine && j + 3 < i));
                }
                for (; j < b0 + sz; ++j) {
                    const uint32_t lj = lo32[4 * j];
                    rank += (lj < mine || (lj == mine && j < i)) ? 1u : 0u;
                }
            } else {
                for (uint32_t j = b0; j < b0 + sz; ++j) {
                    const unsigned long long lj = sm.buf[0][j].low;
                    rank += (lj < it.low || (lj == it.low && j < i)) ? 1u : 0u;
                }
            }
            sm.buf[1][b0 + rank] = it;
        }
        __syncthreads();
        cur = 1;
    }
    // ---- otherwise: stable LSD passes in shared memory (ranking as in radix_pass_kernel: ballot-built
    // peer masks); every warp takes an equal share of the items, a multiple of 32
    const uint32_t blk = ((m + UNIT_WARPS * 32 - 1) / (UNIT_WARPS * 32)) * 32, rounds = blk / 32;
    for (uint32_t pass = 0; pass < (small_bins ? 0u : npass); ++pass) {
        for (uint32_t i = tid; i < UNIT_WARPS * RADIX_BINS; i += UNIT_THREADS) (&sm.whist[0])[i] = 0;
        __syncthreads();
        MsdItem it[UNIT_ITEMS];
        uint32_t dg[UNIT_ITEMS], rk[UNIT_ITEMS];
        const uint32_t wbase = warp * blk;
#pragma unroll
        for (int i = 0; i < UNIT_ITEMS; ++i) {
            if ((uint32_t)i >= rounds) break;          // warp-uniform
            const uint32_t idx = wbase + i * 32 + lane;
            const bool valid = idx < m;
            if (valid) it[i] = sm.buf[cur][idx];
            else { it[i].low = 0; it[i].val = 0; it[i].bin = 0; }
            const uint32_t d = msd_pass_digit(it[i].low, it[i].bin & ~MSD_FLAG8, low_bits, pass);
            dg[i] = d;
            const uint32_t act = __ballot_sync(0xFFFFFFFFu, valid);
            uint32_t peers = act;
#pragma unroll
            for (int bit = 0; bit < RADIX_BITS; ++bit) {
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, (d >> bit) & 1u);
                peers &= ((d >> bit) & 1u) ? bal : ~bal;
            }
            uint32_t r = 0;
            if (valid) {
                const uint32_t leader = __ffs((int)peers) - 1;
                uint32_t old = 0;
                if (lane == leader) { old = sm.whist[warp * RADIX_BINS + d]; sm.whist[warp * RADIX_BINS + d] = old + __popc(peers); }
                old = __shfl_sync(act, old, leader);
                r = old + __popc(peers & lt);
            }
            rk[i] = r;
            __syncwarp();
        }
        __syncthreads();
        unsigned int tcount = 0, tincl = 0;
        if (tid < RADIX_BINS) {
            unsigned int run = 0;
#pragma unroll
            for (int w = 0; w < UNIT_WARPS; ++w) { const unsigned int c = sm.whist[w * RADIX_BINS + tid]; sm.whist[w * RADIX_BINS + tid] = run; run += c; }
            tcount = run;
            tincl = run;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { unsigned int v = __shfl_up_sync(0xFFFFFFFFu, tincl, d); if (lane >= d) tincl += v; }
            if (lane == 31) sm.scan_tmp[warp] = tincl;
        }
        __syncthreads();
        if (tid < RADIX_BINS) {
            unsigned int wsum = 0;
            for (uint32_t w = 0; w < warp; ++w) wsum += sm.scan_tmp[w];
            sm.bin_excl[tid] = wsum + tincl - tcount;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < UNIT_ITEMS; ++i) {
            if ((uint32_t)i >= rounds) break;
            const uint32_t idx = wbase + i * 32 + lane;
            if (idx < m) sm.buf[cur ^ 1][sm.bin_excl[dg[i]] + sm.whist[warp * RADIX_BINS + dg[i]] + rk[i]] = it[i];
        }
        __syncthreads();
        cur ^= 1;
    }
    MsdItem *sorted = sm.buf[cur];
    UT(3);
    ulonglong2 *skey = (ulonglong2 *)sm.buf[cur ^ 1];       // (k0, k1) of sorted[i], filled below
    unsigned long long *scnt = (unsigned long long *)&sm.whist[0];

    auto same_word = [&](uint32_t a, uint32_t b) { return sorted[a].low == sorted[b].low && sorted[a].bin == sorted[b].bin; };
    // ---- entries of the sorted items: one 32-byte gather each; key, value and partition stay in shared
    // memory at the item's sorted position (the entry index has done its work: its slot takes the partition)
    const uint32_t base = tid * UNIT_ITEMS;
    for (uint32_t i = tid; i < m; i += UNIT_THREADS) {
        const ulonglong4 e = ld256(&ent.e[sorted[i].val]);
        skey[i] = make_ulonglong2(e.x, e.y);
        scnt[i] = e.z;
        sorted[i].val = (uint32_t)e.w;
    }
    __syncthreads();
    UT(4);
    // ---- keys that agree in the whole sort word and have 8+ bytes are ordered by their later bytes.
    // First every item says (iflag) whether it continues its predecessor's sort word (1) and, if so,
    // whether it is another key (2); only groups that hold different keys are ordered.  Copies of one
    // key, which is what a shuffle delivers most, cost one 16-byte compare each.
    unsigned char *iflag = sm.iflag;
    auto same_key_at = [&](uint32_t x, uint32_t y) {
        const ulonglong2 ka = skey[x], kb = skey[y];
        if (ka.x == kb.x && ka.y == kb.y) return true;
        if ((ka.y >> 56) != 0xFFu) return false;          // inline keys: the 128 bits ARE the key
        Key A, B; A.k0 = ka.x; A.k1 = ka.y; B.k0 = kb.x; B.k1 = kb.y;
        return same_key(A, B, bytes);
    };
    for (uint32_t i = tid; i < m; i += UNIT_THREADS) {
        uint32_t f = 0;
        if (i > 0 && same_word(i, i - 1)) {
            f = 1;
            if ((sorted[i].bin & MSD_FLAG8) && !same_key_at(i, i - 1)) f = 3;
        }
        iflag[i] = (unsigned char)f;
    }
    __syncthreads();
    for (uint32_t i = tid; i < m; i += UNIT_THREADS) {       // first members: does the group hold different keys?
        if ((iflag[i] & 1) || !(sorted[i].bin & MSD_FLAG8)) continue;
        uint32_t g = 1, need = 0;
        while (i + g < m && g <= MSD_TIE_GROUP && (iflag[i + g] & 1)) { need |= iflag[i + g] & 2u; ++g; }
        if (g > MSD_TIE_GROUP) { atomicOr(&ctl->msd_fallback, MSD_FB_TIES); sm.fallback = 1; continue; }
        if (!need) continue;
        for (uint32_t a2 = 0; a2 < g; ++a2) iflag[i + a2] |= 8u;           // the members order themselves below
        atomicAdd(&ctl->tie_small, 1u);
    }
    __syncthreads();
    // Every member of such a group counts the members that come before it in the full order (g whole-key
    // compares, all members at once -- not g^2 / 2 by one thread), takes key, value and partition into
    // registers, and after the barrier stores them at group begin + count.
    {
        ulonglong2 mk[UNIT_ITEMS];
        unsigned long long mc[UNIT_ITEMS];
        uint32_t mv[UNIT_ITEMS], md[UNIT_ITEMS];
#pragma unroll
        for (int r = 0; r < UNIT_ITEMS; ++r) {
            const uint32_t i = tid + (uint32_t)r * UNIT_THREADS;
            md[r] = 0xFFFFFFFFu;
            mk[r] = make_ulonglong2(0, 0); mc[r] = 0; mv[r] = 0;
            if (i >= m || !(iflag[i] & 8u)) continue;
            uint32_t s0 = i;
            while (iflag[s0] & 1) --s0;
            uint32_t g = 1;
            while (s0 + g < m && (iflag[s0 + g] & 1)) ++g;              // <= MSD_TIE_GROUP, checked above
            const ulonglong2 kx2 = skey[i];
            Key kx; kx.k0 = kx2.x; kx.k1 = kx2.y;
            // two inline keys (<= 15 bytes, the usual case) that agree in bytes 0..6: bytes 7..14 as one
            // big-endian word (zero-padded), then the length -- the order key_less gives, in ten instructions
            const bool x_inline = !key_is_long(kx);
            uint32_t rank = 0;
            for (uint32_t a2 = 0; a2 < g; ++a2) {
                const uint32_t j = s0 + a2;
                if (j == i) continue;
                Key kb; kb.k0 = skey[j].x; kb.k1 = skey[j].y;
                bool less;                                              // kb before kx; equal keys keep their order
                if (x_inline && !key_is_long(kb)) {
                    less = inline_tail_less(kb, kx, j < i);
                } else {
                    less = key_less(kb, kx, j, i, bytes, 1);
                }
                rank += less ? 1u : 0u;
            }
            if (s0 + rank != i) { md[r] = s0 + rank; mk[r] = kx2; mc[r] = scnt[i]; mv[r] = sorted[i].val; }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < UNIT_ITEMS; ++r)
            if (md[r] != 0xFFFFFFFFu) { skey[md[r]] = mk[r]; scnt[md[r]] = mc[r]; sorted[md[r]].val = mv[r]; }
    }
    __syncthreads();
    // a unit that declined still takes part in the look-back below (with nothing), so that no later
    // unit waits for it; the rows are rewritten by the LSD path anyway
    const bool declined = sm.fallback != 0;
    // heads (after the ordering): an item starts a row unless it repeats its predecessor's key
    for (uint32_t i = tid; i < m; i += UNIT_THREADS) {
        bool head = i == 0 || !same_word(i, i - 1);   // the first item of a unit starts a bin: always a new key
        if (!head && (sorted[i].bin & MSD_FLAG8)) head = !same_key_at(i, i - 1);
        iflag[i] = head ? 4 : 0;
    }
    __syncthreads();
    UT(5);
    Key key[UNIT_ITEMS];
    uint32_t part[UNIT_ITEMS], len[UNIT_ITEMS];
    uint32_t headmask = 0, nh = 0;
    unsigned long long nb = 0;
#pragma unroll
    for (int j = 0; j < UNIT_ITEMS; ++j) {
        const uint32_t i = base + j;
        part[j] = 0; len[j] = 0; key[j].k0 = 0; key[j].k1 = 0;
        if (i < m && !declined) {
            key[j].k0 = skey[i].x; key[j].k1 = skey[i].y;
            part[j] = sorted[i].val;
            len[j] = key_len(key[j]);
            if (iflag[i]) { headmask |= 1u << j; ++nh; nb += len[j]; }
        }
    }
    // merged values (merge_min_keys, utils.lua:232-247, with the job's reducer): a thread folds the runs
    // inside its own items; a run that began in an earlier thread's items gets this thread's part added
    // atomically after the heads have stored theirs.  The sums live in the sorted items' `low` slots,
    // whose sort words are not needed any more.
    const bool additive = op == OP_SUM || op == OP_COUNT;
    auto fold = [&](unsigned long long x, unsigned long long y) { return additive ? x + y : max(x, y); };
    unsigned long long lead_acc = 0, acc = 0;
    bool lead = false;
    int open_j = -1;
#pragma unroll
    for (int j = 0; j < UNIT_ITEMS; ++j) {
        const uint32_t i = base + j;
        if (i >= m || declined) break;
        const unsigned long long v = scnt[i];
        if ((headmask >> j) & 1u) {
            if (open_j >= 0) sorted[base + open_j].low = acc;
            open_j = j; acc = v;
        } else if (open_j >= 0) {
            acc = fold(acc, v);
        } else {
            lead_acc = lead ? fold(lead_acc, v) : v;
            lead = true;
        }
    }
    if (open_j >= 0) sorted[base + open_j].low = acc;
    __syncthreads();
    if (lead) {
        uint32_t h = base - 1;                        // the run's head lies in an earlier thread's items
        while (!iflag[h]) --h;
        if (additive) atomicAdd(&sorted[h].low, lead_acc);
        else atomicMax(&sorted[h].low, lead_acc);
    }
    __syncthreads();
    unsigned long long cval[UNIT_ITEMS];
#pragma unroll
    for (int j = 0; j < UNIT_ITEMS; ++j) cval[j] = ((headmask >> j) & 1u) ? sorted[base + j].low : 0ull;
    uint32_t ih = nh;
    unsigned long long ib = nb;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t vh = __shfl_up_sync(0xFFFFFFFFu, ih, d);
        const unsigned long long vb = __shfl_up_sync(0xFFFFFFFFu, ib, d);
        if (lane >= d) { ih += vh; ib += vb; }
    }
    if (lane == 31) { sm.wh[warp] = ih; sm.wb[warp] = ib; }
    __syncthreads();
    uint32_t oh = ih - nh, th = 0;
    unsigned long long ob = ib - nb, tb = 0;
#pragma unroll
    for (int w = 0; w < UNIT_WARPS; ++w) {
        if (w < (int)warp) { oh += sm.wh[w]; ob += sm.wb[w]; }
        th += sm.wh[w]; tb += sm.wb[w];
    }
    // the unit's key bytes are laid out in shared memory now (offsets inside the unit are known; where the
    // unit starts in the output is not yet), so that this work overlaps the wait for the units before
    // (staged in the buffer that held the keys: every thread has its keys in registers since the barrier
    // after the heads' stores.  Four units of 1024 / three of 1536 items then fit the SM's 196 KB
    // shared-memory configuration -- the 228 KB one slows shared-memory accesses, see map_kernel.cuh.)
    uint8_t *stage = (uint8_t *)skey;
    constexpr uint32_t UNIT_STAGE = UNIT_CAP * 16 - 16;   // more key bytes than that: written directly
    const bool staged = !declined && tb <= UNIT_STAGE;
    if (staged) {
        unsigned long long lb = ob;
#pragma unroll
        for (int j = 0; j < UNIT_ITEMS; ++j)
            if ((headmask >> j) & 1u) { write_key_bytes(stage + (uint32_t)lb, key[j], len[j], bytes); lb += len[j]; }
    }
    UT(6);
    // jobs after a shuffle (the same key may have arrived from several ranks): the unit knows its rows
    // only now, after sorting and merging
    const unsigned long long my_last = (m && !declined) ? (unsigned long long)(sorted[m - 1].val + 1u) : 0ull;
    if (!early && warp == 0) lookback(th, tb, my_last);
    __syncthreads();
    if (declined) return;
    UT(7);
    unsigned long long o = sm.prefix_h + oh, boff = sm.prefix_b + ob;
    const unsigned long long unit_b0 = sm.prefix_b;
    const bool overflow = unit_b0 + tb > key_bytes_cap;
    if (overflow && tid == 0) atomicOr(&ctl->err, ERR_KEYS_FULL);
    const bool last_unit = unit == nunits - 1;
#pragma unroll
    for (int j = 0; j < UNIT_ITEMS; ++j) {
        const uint32_t i = base + j;
        if (i >= m) break;
        if (!((headmask >> j) & 1u)) continue;
        rows.part[o] = part[j];
        rows.count[o] = cval[j];
        rows.key_off[o] = boff;
        if (!staged && !overflow) write_key_bytes(rows.key_bytes + boff, key[j], len[j], bytes);
        // partition starts: rows are partition-sorted; the row before this one is the previous head
        // of the unit or, for the unit's first row, the last row of the units before (look-back)
        uint32_t p_first;
        bool new_part;
        if (i == 0) {
            p_first = sm.prefix_part;                   // (last partition before) + 1, or 0 if there is no row before
            new_part = sm.prefix_part == 0 || sm.prefix_part - 1 != part[j];
        } else {
            const uint32_t pp = j > 0 ? part[j - 1] : sorted[i - 1].val;
            p_first = pp + 1;
            new_part = pp != part[j];
        }
        if (new_part)
            for (uint32_t q = p_first; q <= part[j] && q <= num_reducers; ++q) rows.part_begin[q] = o;
        boff += len[j];
        ++o;
    }
    // totals and sentinels: the thread that holds the last item of the last unit
    if (last_unit && m && base < m && base + UNIT_ITEMS >= m) {
        ctl->n_out = o;
        ctl->key_bytes = boff;
        rows.key_off[o] = boff;
        for (uint32_t q = sorted[m - 1].val + 1u; q <= num_reducers; ++q) rows.part_begin[q] = o;
    }
    if (last_unit && m == 0 && tid == 0) {
        // cannot happen (the last unit holds the last bin start), kept for safety
        ctl->n_out = sm.prefix_h; ctl->key_bytes = sm.prefix_b; rows.key_off[sm.prefix_h] = sm.prefix_b;
    }
    if (staged && !overflow) {
        // the staged bytes start at the unit's first output byte, whose alignment is only known now:
        // whole 16-byte chunks of the OUTPUT are assembled from five staged words and one funnel shift each
        UT(8);
        const uint32_t mis = (uint32_t)(unit_b0 & 15);
        uint8_t *dst = rows.key_bytes + (unit_b0 - mis);                       // 16-byte aligned
        const uint32_t end = mis + (uint32_t)tb;                               // in dst coordinates
        const uint32_t *sw = (const uint32_t *)stage;
        for (uint32_t cb = tid * 16; cb < end; cb += UNIT_THREADS * 16) {
            if (cb >= mis && cb + 16 <= end) {
                const uint32_t so = cb - mis, wi = so >> 2, sh = (so & 3) * 8;
                const uint32_t x0 = sw[wi], x1 = sw[wi + 1], x2 = sw[wi + 2], x3 = sw[wi + 3], x4 = sw[wi + 4];
                *(uint4 *)(dst + cb) = make_uint4(funnel_r(x0, x1, sh), funnel_r(x1, x2, sh), funnel_r(x2, x3, sh), funnel_r(x3, x4, sh));
            } else {
                for (uint32_t q = cb < mis ? mis : cb; q < cb + 16 && q < end; ++q) dst[q] = stage[q - mis];
            }
        }
    }
    UT(9);
}

} // namespace mr
