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
// K6 and K7 are one single-pass kernel (decoupled look-back over tiles).
#pragma once
#include "common.cuh"

namespace mr {

constexpr int RADIX_BITS = 8;
constexpr int RADIX_BINS = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 12;          // 96 bits / 8
constexpr int SORT_THREADS = 512;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // 4096
constexpr int LOOKBACK = 32;
constexpr int ROWS_THREADS = 256;
constexpr int ROWS_ITEMS = 8;
constexpr int ROWS_TILE = ROWS_THREADS * ROWS_ITEMS; // 2048
constexpr uint32_t ROWS_STAGE = ROWS_TILE * INLINE_MAX; // key bytes of a tile of inline keys
constexpr uint32_t TIE_SMALL_GROUP = 16;  // tie groups up to this size are ordered by one thread
constexpr uint32_t TIE_SMALL_MAX = 1024;  // members of larger groups one CTA can order

struct DigitRun { uint8_t shift, width, dst, pad; }; // shift 0..63 -> lo, 64..95 -> hi

// what the radix sort moves: 96-bit sort key + index of the entry
struct __align__(16) SortItem {
    unsigned long long lo; // 7 key bytes big-endian << 8 | min(len, 8)
    uint32_t hi;           // partition
    uint32_t val;          // entry index
};

// Plan of the most-significant-digit path (msd_kernels.cuh).  The sort word is read as 9 fields,
// most significant first: partition, key bytes 0..6, min(len, 8).  A key byte is replaced by its RANK
// among the byte values that occur at that position in this job (Ctl::occ), so the codes are dense
// whatever the alphabet is; the bin of an item is the mixed-radix number of its first k fields times
// e, plus a monotone e-way split of field k; the rest of the order is the bit-packed ranks of the
// fields k..8 ("low").
constexpr int MSD_FIELDS = 9;
struct MsdPlan {
    uint32_t nbins, k, e, low_bits, fallback, part_base;   // part_base: the smallest partition present (field 0 counts from it)
    uint32_t card[MSD_FIELDS];      // values a field can take (partition: num_reducers; bytes: occurring values; length: 9)
    uint8_t width[MSD_FIELDS];      // bits of a field's rank in the packed low word
    uint8_t pad[3];
    uint32_t pad2;
};

struct Ctl {
    unsigned long long n;            // entries (distinct keys before the merge)
    unsigned long long or_lo, and_lo;
    unsigned int or_hi, and_hi;
    unsigned int part_min, part_max; // smallest / largest partition among the sort items (starts at ~0 / 0), for the MSD plan
    unsigned int occ[7][8];          // byte values that occur at key byte 0..6 (256-bit masks), for the MSD plan
    unsigned int npasses;
    unsigned int err;
    unsigned int n_tie;              // members of tie groups larger than TIE_SMALL_GROUP
    unsigned int tie_overflow;       // n_tie > TIE_SMALL_MAX: host must run the general path
    unsigned int tie_small;          // tie groups ordered in place
    unsigned int rows_ticket;
    unsigned long long n_out;        // rows after the segmented reduce
    unsigned long long key_bytes;    // sum of key lengths over rows
    unsigned long long tokens;
    unsigned long long pool_used;    // bytes of the key pool in use (bigram keys longer than 15 bytes)
    unsigned int max_tie_len;        // longest key among the members of large tie groups
    unsigned int reserved0;
    unsigned long long sent;         // peer-to-peer shuffle: records this rank sent (its map_keys)
    unsigned int nruns[MAX_PASSES];
    DigitRun runs[MAX_PASSES][8];
    unsigned int tickets[MAX_PASSES];
    unsigned int ghist[MAX_PASSES][RADIX_BINS];
    long long dbg[16];               // MRGPU_SORT_TIMING builds: clock64 stamps of one tile
    MsdPlan msd_plan;                // msd_kernels.cuh
    unsigned int msd_fallback;       // non-zero: the MSD path declined this job (MSD_FB_* bits)
    unsigned int msd_pad;
    unsigned char msd_lut[7][256];   // rank of a byte value at key byte 0..6
    unsigned char msd_lutk[256];     // the e-way split of the plan's field k, when that is a key byte
};

// Dense entries: the occupied slots of the global table copied side by side, 32 bytes each
// {k0, k1, count, partition}, so that one 32-byte gather fetches everything about a key.
struct EntryArrays {
    GEntry *e;
    __device__ __forceinline__ Key key(uint64_t i) const
    {
        ulonglong2 v = *(const ulonglong2 *)&e[i];
        Key k; k.k0 = v.x; k.k1 = v.y;
        return k;
    }
    __device__ __forceinline__ unsigned long long cnt(uint64_t i) const { return e[i].cnt; }
    __device__ __forceinline__ uint32_t part(uint64_t i) const { return (uint32_t)e[i].pad; }
    __device__ __forceinline__ uint32_t owner(uint64_t i) const { return (uint32_t)(e[i].pad >> 32); } // set by compact_partition_kernel
};

// one 256-bit load (sm_100: ld.global.v4.u64), e.g. a whole table entry = one 32-byte sector
__device__ __forceinline__ ulonglong4 ld256(const void *p)
{
    ulonglong4 r;
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
                 : "l"(p));
    return r;
}

// 32-byte load that takes part in the memory model: for data this kernel also writes (table slots
// cleared after they were read) or that a peer stored while the kernel was already running
__device__ __forceinline__ ulonglong4 ld256_coherent(const void *p)
{
    ulonglong4 r;
    asm volatile("ld.global.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
                 : "l"(p) : "memory");
    return r;
}

// ------------------------------------------------------------------ K3
// which byte values occur at each of the 7 key-byte positions of the sort word: per-CTA masks in
// shared memory (zeroed by occ_clear before the first occ_add), merged into Ctl::occ by orand_commit
__device__ __forceinline__ void occ_clear(unsigned int *s_occ)
{
    if (threadIdx.x < 56) s_occ[threadIdx.x] = 0;
    __syncthreads();
}
__device__ __forceinline__ void occ_add(unsigned int *s_occ, unsigned long long chunk)
{
#pragma unroll
    for (int p = 0; p < 7; ++p) {
        const uint32_t b = (uint32_t)(chunk >> (56 - 8 * p)) & 0xFFu;
        const uint32_t bit = 1u << (b & 31);
        unsigned int *w = &s_occ[p * 8 + (b >> 5)];
        if (!(*(volatile unsigned int *)w & bit)) atomicOr(w, bit);
    }
}

__device__ __forceinline__ void orand_commit(Ctl *ctl, uint32_t hi_or, uint32_t hi_and,
                                             uint64_t lo_or, uint64_t lo_and, bool any, const unsigned int *s_occ = nullptr,
                                             uint32_t pmin = ~0u, uint32_t pmax = 0)
{
    if (s_occ) {
        __syncthreads();
        if (threadIdx.x < 56 && s_occ[threadIdx.x]) atomicOr(&ctl->occ[0][0] + threadIdx.x, s_occ[threadIdx.x]);
        for (int d = 16; d; d >>= 1) {
            pmin = min(pmin, __shfl_xor_sync(0xFFFFFFFFu, pmin, d));
            pmax = max(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, d));
        }
        if ((threadIdx.x & 31) == 0 && pmin <= pmax) { atomicMin(&ctl->part_min, pmin); atomicMax(&ctl->part_max, pmax); }
    }
    // warp-reduce, block-reduce, then one set of atomics per CTA (same-address atomics serialise)
    __shared__ uint32_t s_hi_or[8], s_hi_and[8], s_any[8];
    __shared__ unsigned long long s_lo_or[8], s_lo_and[8];
    for (int d = 16; d; d >>= 1) {
        uint32_t o_hi_or = __shfl_xor_sync(0xFFFFFFFFu, hi_or, d);
        uint32_t o_hi_and = __shfl_xor_sync(0xFFFFFFFFu, hi_and, d);
        uint64_t o_lo_or = __shfl_xor_sync(0xFFFFFFFFu, lo_or, d);
        uint64_t o_lo_and = __shfl_xor_sync(0xFFFFFFFFu, lo_and, d);
        hi_or |= o_hi_or; hi_and &= o_hi_and; lo_or |= o_lo_or; lo_and &= o_lo_and;
    }
    uint32_t anym = __ballot_sync(0xFFFFFFFFu, any);
    const uint32_t warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        s_hi_or[warp] = hi_or; s_hi_and[warp] = hi_and; s_lo_or[warp] = lo_or; s_lo_and[warp] = lo_and;
        s_any[warp] = anym;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0;
        for (uint32_t w = 0; w < blockDim.x / 32; ++w) {
            a |= s_any[w];
            if (!s_any[w]) continue;
            hi_or |= s_hi_or[w]; hi_and &= s_hi_and[w]; lo_or |= s_lo_or[w]; lo_and &= s_lo_and[w];
        }
        if (a) {
            atomicOr(&ctl->or_hi, hi_or);
            atomicAnd(&ctl->and_hi, hi_and);
            atomicOr(&ctl->or_lo, (unsigned long long)lo_or);
            atomicAnd(&ctl->and_lo, (unsigned long long)lo_and);
        }
    }
}

// One atomic per 1024-slot tile reserves the output range (a per-warp atomic on one address
// costs ~1.4 ns each).
constexpr int COMPACT_ITEMS = 4;

// table slots -> dense entries WITH their partition ids (and owner ranks), sort items and the
// varying-bit masks, in one read of the table.  A tile's occupied slots are first packed side by
// side in shared memory, so that the partitioner (a byte loop per key) runs on full warps
// although only a fraction of the slots is occupied.
__global__ void __launch_bounds__(256) compact_partition_kernel(
    GEntry *gt, uint64_t nslots, const uint8_t *bytes, uint32_t num_reducers, uint32_t partition_kind,
    EntryArrays ent, SortItem *items, uint64_t max_keys, Ctl *ctl, OwnerMap owners, uint32_t *lens)
{
    constexpr int ITEMS = COMPACT_ITEMS;
    __shared__ ulonglong4 s_ent[256 * ITEMS];
    __shared__ unsigned int wsum[8];
    __shared__ unsigned long long s_base;
    __shared__ unsigned int s_occ[56];
    occ_clear(s_occ);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t tile_slots = 256 * ITEMS;
    const uint64_t ntiles = (nslots + tile_slots - 1) / tile_slots;
    uint32_t hi_or = 0, hi_and = ~0u, pmin = ~0u, pmax = 0;
    uint64_t lo_or = 0, lo_and = ~0ull;
    bool any = false;
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        ulonglong4 e[ITEMS];
        uint32_t nocc = 0;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint64_t slot = tile * tile_slots + (uint64_t)i * 256 + threadIdx.x;
            e[i] = make_ulonglong4(0, 0, 0, 0);
            if (slot < nslots) e[i] = ld256_coherent(&gt[slot]);
            if (e[i].y != 0) {
                // leave the table empty for the next job: only occupied slots are written, which is
                // a fraction of what clearing the whole table costs
                ((ulonglong4 *)gt)[slot] = make_ulonglong4(0, 0, 0, 0);
                ++nocc;
            }
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
        uint32_t pos = woff + incl - nocc;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (e[i].y != 0) s_ent[pos++] = e[i];
        __syncthreads();
        const unsigned long long base = total ? s_base : 0;
        for (uint32_t t = threadIdx.x; t < total; t += 256) {
            const uint64_t idx = base + t;
            if (idx >= max_keys) { atomicOr(&ctl->err, ERR_KEYS_FULL); continue; }
            ulonglong4 v = s_ent[t];
            Key key; key.k0 = v.x; key.k1 = v.y;
            const uint32_t part = partition_kind == 1 ? __byte_perm((uint32_t)key.k0, 0, 0x0123) % num_reducers
                                                      : partition_of_key(key, bytes, num_reducers);
            const uint64_t chunk = sort_chunk(key, bytes, 0);
            const uint32_t owner = owners.world > 1 ? owner_of(owners, part, chunk) : 0;
            v.w = part | (unsigned long long)owner << 32; // EntryArrays::part / ::owner
            ((ulonglong4 *)ent.e)[idx] = v;
            SortItem it; it.lo = chunk; it.hi = part; it.val = (uint32_t)idx;
            items[idx] = it;
            lens[idx] = key_len(key);              // for the MSD path of jobs without a shuffle (msd_kernels.cuh)
            hi_or |= part; hi_and &= part; lo_or |= chunk; lo_and &= chunk;
            pmin = min(pmin, part); pmax = max(pmax, part);
            occ_add(s_occ, chunk);
            any = true;
        }
        __syncthreads();
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any, s_occ, pmin, pmax);
}

// dense entries -> partition id + sort item (+ which key bits vary at all)
// partition_kind 0: the WordCount partitioner (Lua 5.2 double FNV, partitionfn.lua:8-16);
// partition_kind 1: `key % NUM_REDUCERS` on a number key carried as 4 big-endian bytes
__global__ void __launch_bounds__(256) partition_entries_kernel(
    const uint8_t *bytes, uint32_t num_reducers, uint32_t partition_kind, EntryArrays ent, SortItem *items,
    uint64_t max_keys, Ctl *ctl)
{
    __shared__ unsigned int s_occ[56];
    occ_clear(s_occ);
    uint32_t hi_or = 0, hi_and = ~0u, pmin = ~0u, pmax = 0;
    uint64_t lo_or = 0, lo_and = ~0ull;
    bool any = false;
    const uint64_t n = ctl->n < max_keys ? ctl->n : max_keys;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        Key key = ent.key(idx);
        uint32_t part = partition_kind == 1 ? __byte_perm((uint32_t)key.k0, 0, 0x0123) % num_reducers
                                            : partition_of_key(key, bytes, num_reducers);
        uint64_t chunk = sort_chunk(key, bytes, 0);
        ent.e[idx].pad = part;
        SortItem it; it.lo = chunk; it.hi = part; it.val = (uint32_t)idx;
        items[idx] = it;
        hi_or |= part; hi_and &= part; lo_or |= chunk; lo_and &= chunk;
        pmin = min(pmin, part); pmax = max(pmax, part);
        occ_add(s_occ, chunk);
        any = true;
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any, s_occ, pmin, pmax);
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

__host__ __device__ __forceinline__ uint32_t extract_digit(const DigitRun *runs, uint32_t nruns, uint32_t hi,
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

// The digit plan again, run by run instead of bit by bit, into whatever arrays the caller has
// (shared memory in radix_hist_kernel).  Same plan as sort_setup_kernel.
__host__ __device__ inline uint32_t build_digit_plan(uint64_t vlo, uint32_t vhi, DigitRun (*runs)[8], unsigned int *nruns)
{
    uint32_t pass = 0, used = 0, nr = 0;
    for (int half = 0; half < 2; ++half) {
        uint64_t v = half ? (uint64_t)vhi : vlo; // runs do not straddle the lo/hi boundary
        const uint32_t base = half ? 64 : 0;
        while (v) {
            const uint32_t start = ffs64(v) - 1;
            const uint64_t ones = ~(v >> start);           // zero bits of the shifted word
            uint32_t w = ones ? ffs64(ones) - 1 : 64 - start;
            if (w > RADIX_BITS - used) w = RADIX_BITS - used;
            DigitRun r; r.shift = (uint8_t)(base + start); r.width = (uint8_t)w; r.dst = (uint8_t)used; r.pad = 0;
            runs[pass][nr++] = r;
            used += w;
            v &= ~((w >= 64 ? ~0ull : ((1ull << w) - 1)) << start);
            if (used == RADIX_BITS) { nruns[pass] = nr; ++pass; used = 0; nr = 0; }
        }
    }
    if (used) { nruns[pass] = nr; ++pass; }
    return pass;
}

// digit histograms of every pass in one read of the keys.  Every CTA derives the digit plan from
// the varying-bit masks for itself (a few dozen instructions of one thread, hidden behind the
// zeroing of the histogram); CTA 0 publishes it for the passes and clamps n.
__global__ void __launch_bounds__(256) radix_hist_kernel(const SortItem *items, unsigned long long max_keys, Ctl *ctl,
                                                         uint32_t *status, uint32_t max_tiles)
{
    __shared__ unsigned int h[MAX_PASSES][RADIX_BINS];
    __shared__ DigitRun runs[MAX_PASSES][8];
    __shared__ unsigned int nruns[MAX_PASSES];
    __shared__ unsigned int s_np;
    unsigned long long n = ctl->n;
    const bool over = n > max_keys;
    if (over) n = max_keys;
    if (blockIdx.x && (uint64_t)blockIdx.x * blockDim.x >= n) return;
    if (threadIdx.x == 0) {
        uint64_t vlo = ctl->or_lo ^ ctl->and_lo;
        uint32_t vhi = ctl->or_hi ^ ctl->and_hi;
        if (n < 2) { vlo = 0; vhi = 0; }
        s_np = build_digit_plan(vlo, vhi, runs, nruns);
    }
    for (uint32_t i = threadIdx.x; i < MAX_PASSES * RADIX_BINS; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t np = s_np;
    if (blockIdx.x == 0) {
        if (threadIdx.x < np * 8) (&ctl->runs[0][0])[threadIdx.x] = (&runs[0][0])[threadIdx.x];
        if (threadIdx.x < np) ctl->nruns[threadIdx.x] = nruns[threadIdx.x];
        if (threadIdx.x == 0) {
            ctl->npasses = np;
            if (over) { ctl->n = max_keys; atomicOr(&ctl->err, ERR_KEYS_FULL); } // reported by the host
        }
    }
    if (np == 0) return;
    {
        // the look-back status words of the passes to come (only the passes and tiles that exist)
        uint64_t nact = (n + blockDim.x - 1) / blockDim.x;
        if (nact < 1) nact = 1;
        if (nact > gridDim.x) nact = gridDim.x;
        const uint64_t ntiles = (n + SORT_TILE - 1) / SORT_TILE, per_pass = ntiles * RADIX_BINS;
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < np * per_pass; i += nact * blockDim.x)
            status[(i / per_pass) * max_tiles * RADIX_BINS + i % per_pass] = 0;
    }
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const SortItem it = items[i];
        for (uint32_t p = 0; p < np; ++p) atomicAdd(&h[p][extract_digit(runs[p], nruns[p], it.hi, it.lo)], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < np * RADIX_BINS; i += blockDim.x) {
        unsigned int c = (&h[0][0])[i];
        if (c) atomicAdd(&(&ctl->ghist[0][0])[i], c);
    }
}

// tile status word: bits 30..31 = state (0 empty, 1 aggregate, 2 inclusive prefix)
constexpr uint32_t ST_AGG = 1u << 30, ST_INC = 2u << 30, ST_VAL = (1u << 30) - 1;
// a look-back that polls this often without progress gives up (about 10 s of L2 round trips): the
// job then fails with an error instead of hanging the stream
constexpr uint32_t LOOKBACK_SPIN_LIMIT = 1u << 24;
constexpr uint32_t ERR_LOOKBACK_TIMEOUT = 32u; // joins the ERR_* bits of common.cuh (16 = ERR_PEER_TIMEOUT)

struct SortBufs {
    SortItem *item[2];
};

// One LSD pass (stable).  Per tile of 4096 items: digits, warp-local ranks from ballot-built
// peer masks, per-bin tile counts published for a decoupled look-back, items reordered by digit in
// shared memory so that the global scatter writes runs of consecutive 16-byte items instead of one
// 32-byte sector per item.  The tile id is an atomic ticket (Ctl::tickets[pass]), not the block
// index: a tile only ever waits for tiles whose CTA has already started, whatever order the
// hardware dispatches blocks in; the wait itself is bounded (LOOKBACK_SPIN_LIMIT polls, seconds)
// and ends in ERR_LOOKBACK_TIMEOUT instead of a hang.
struct __align__(16) SortSmem {
    SortItem sorted[SORT_TILE];                 // 64 KB: the tile in digit order
    unsigned int whist[SORT_WARPS][RADIX_BINS]; // per-warp digit counts -> exclusive offsets
    unsigned int bin_base[RADIX_BINS];          // global start of the tile's items of each digit
    unsigned int tile_excl[RADIX_BINS];         // start of each digit inside the tile
    unsigned int scan_tmp[2][RADIX_BINS / 32];
    unsigned char dig[SORT_TILE];               // digit of sorted[i]
    DigitRun runs[8];
    unsigned int nruns;
    unsigned int tile;                          // this CTA's ticket
};

#ifdef MRGPU_SORT_TIMING
#define STAMP(k) do { if (blockIdx.x == 64 && pass == 1 && threadIdx.x == 0) ctl->dbg[k] = clock64(); } while (0)
#else
#define STAMP(k) do { } while (0)
#endif

__global__ void __launch_bounds__(SORT_THREADS) radix_pass_kernel(SortBufs bufs, uint32_t pass,
                                                                  uint32_t *status, uint32_t max_tiles,
                                                                  Ctl *ctl)
{
    extern __shared__ __align__(16) uint8_t sort_smem_raw[];
    SortSmem &sm = *(SortSmem *)sort_smem_raw;
    STAMP(0);
    if (pass >= ctl->npasses) return;
    const uint64_t n = ctl->n;
    if (threadIdx.x == 0) sm.tile = atomicAdd(&ctl->tickets[pass], 1u);
    __syncthreads();
    const uint32_t tile = sm.tile;
    if ((uint64_t)tile * SORT_TILE >= n) return;
    const SortItem *src = bufs.item[pass & 1];
    SortItem *dst = bufs.item[(pass & 1) ^ 1];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lt = lanemask_lt();

    // item loads first: their latency overlaps the set-up below
    SortItem it[SORT_ITEMS];
    const uint64_t wbase = (uint64_t)tile * SORT_TILE + (uint64_t)warp * (32 * SORT_ITEMS);
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        if (idx < n) it[i] = src[idx];
        else { it[i].lo = 0; it[i].hi = 0; it[i].val = 0; }
    }
    if (threadIdx.x < 8) sm.runs[threadIdx.x] = ctl->runs[pass][threadIdx.x];
    if (threadIdx.x == 8) sm.nruns = ctl->nruns[pass];
    for (uint32_t i = threadIdx.x; i < SORT_WARPS * RADIX_BINS; i += SORT_THREADS) (&sm.whist[0][0])[i] = 0;
    __syncthreads();
    STAMP(1);
    const uint32_t nruns = sm.nruns;

    uint16_t k_rank[SORT_ITEMS];
    uint8_t k_dig[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) k_dig[i] = (uint8_t)extract_digit(sm.runs, nruns, it[i].hi, it[i].lo);
    STAMP(2);
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const bool valid = wbase + (uint64_t)i * 32 + lane < n;
        const uint32_t act = __ballot_sync(0xFFFFFFFFu, valid);
        const uint32_t d = k_dig[i];
        // lanes holding the same digit: eight ballots (one per digit bit), all independent
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
            if (lane == leader) { old = sm.whist[warp][d]; sm.whist[warp][d] = old + __popc(peers); }
            old = __shfl_sync(act, old, leader); // one shuffle for the warp; each lane reads its leader
            r = old + __popc(peers & lt);
        }
        k_rank[i] = (uint16_t)r;
        __syncwarp();
    }
    STAMP(3);
    __syncthreads();
    STAMP(4);
    // thread b < 256 owns bin b: per-warp exclusive offsets, tile count, look-back
    unsigned int prefix = 0, g = 0, incl = 0, tcount = 0, tincl = 0;
    if (threadIdx.x < RADIX_BINS) {
        const uint32_t b = threadIdx.x;
        unsigned int c[SORT_WARPS];
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) c[w] = sm.whist[w][b];
        unsigned int run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) { sm.whist[w][b] = run; run += c[w]; }
        tcount = run;
        // the status word carries its own flag, so no fence is needed around it
        volatile uint32_t *st = status + ((uint64_t)pass * max_tiles) * RADIX_BINS;
        st[(uint64_t)tile * RADIX_BINS + b] = (tile == 0 ? ST_INC : ST_AGG) | tcount;
        g = ctl->ghist[pass][b];
        STAMP(5);
        if (tile > 0) {
            // look back LOOKBACK tiles at a time: the loads are independent, so the inclusive
            // prefix crosses the (co-resident) tiles in tiles / LOOKBACK serial L2 round trips
            int t = (int)tile - 1;
            bool done = false;
            uint32_t polls = 0;
            while (!done) {
                if (++polls > LOOKBACK_SPIN_LIMIT) { atomicOr(&ctl->err, ERR_LOOKBACK_TIMEOUT); break; }
                uint32_t sv[LOOKBACK];
#pragma unroll
                for (int k = 0; k < LOOKBACK; ++k) sv[k] = t - k >= 0 ? st[(uint64_t)(t - k) * RADIX_BINS + b] : (2u << 30);
                int k = 0;
#pragma unroll
                for (; k < LOOKBACK; ++k) {
                    if ((sv[k] >> 30) == 0) break;           // not published yet: re-read from here
                    prefix += sv[k] & ST_VAL;
                    if ((sv[k] >> 30) == 2) { done = true; break; }
                }
                t -= k;
            }
            st[(uint64_t)tile * RADIX_BINS + b] = ST_INC | (prefix + tcount);
        }
        STAMP(6);
        // two exclusive scans over the bins: global histogram (bin starts) and tile counts
        incl = g; tincl = tcount;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, d), tv = __shfl_up_sync(0xFFFFFFFFu, tincl, d);
            if (lane >= d) { incl += v; tincl += tv; }
        }
        if (lane == 31) { sm.scan_tmp[0][warp] = incl; sm.scan_tmp[1][warp] = tincl; }
    }
    __syncthreads();
    if (threadIdx.x < RADIX_BINS) {
        unsigned int wsum = 0, twsum = 0;
        for (uint32_t w = 0; w < warp; ++w) { wsum += sm.scan_tmp[0][w]; twsum += sm.scan_tmp[1][w]; }
        sm.bin_base[threadIdx.x] = wsum + incl - g + prefix;
        sm.tile_excl[threadIdx.x] = twsum + tincl - tcount;
    }
    __syncthreads();
    STAMP(7);
    // reorder by digit in shared memory, then write runs of consecutive items
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        if (wbase + (uint64_t)i * 32 + lane < n) {
            const uint32_t d = k_dig[i];
            const uint32_t q = sm.tile_excl[d] + sm.whist[warp][d] + k_rank[i];
            sm.sorted[q] = it[i];
            sm.dig[q] = (unsigned char)d;
        }
    }
    __syncthreads();
    const uint64_t tile_n = n - (uint64_t)tile * SORT_TILE < SORT_TILE ? n - (uint64_t)tile * SORT_TILE : SORT_TILE;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; ++i) {
        const uint32_t q = (uint32_t)i * SORT_THREADS + threadIdx.x;
        if (q < tile_n) {
            const uint32_t d = sm.dig[q];
            dst[(uint64_t)sm.bin_base[d] + (q - sm.tile_excl[d])] = sm.sorted[q];
        }
    }
    STAMP(8);
}

// ------------------------------------------------------------------ ties beyond 7 bytes
// After the radix sort, keys that agree in (partition, first 7 bytes) and have more bytes
// sit next to each other in arbitrary order ("tie group").  Groups of up to
// TIE_SMALL_GROUP keys (the normal case in text) are ordered in place by one thread;
// members of larger groups are collected and ordered by bitonic networks.
struct TieItem { uint32_t pos, val; };

// full order of two keys of the same partition, round by round (7 key bytes per round)
__device__ inline bool key_less(const Key &ka, const Key &kb, uint32_t va, uint32_t vb, const uint8_t *bytes,
                                uint32_t first_round)
{
    for (uint32_t r = first_round;; ++r) {
        uint64_t ca = sort_chunk(ka, bytes, r), cb = sort_chunk(kb, bytes, r);
        if (ca != cb) return ca < cb;
        if ((ca & 0xFF) < 8) break; // both keys end here: equal keys (same key from two sources)
    }
    return va < vb;
}

__device__ inline bool tie_less(const TieItem &a, const TieItem &b, const EntryArrays &ent,
                                const uint8_t *bytes)
{
    if (a.pos == 0xFFFFFFFFu || b.pos == 0xFFFFFFFFu) return a.pos != 0xFFFFFFFFu && b.pos == 0xFFFFFFFFu;
    uint32_t pa = ent.part(a.val), pb = ent.part(b.val);
    if (pa != pb) return pa < pb;
    return key_less(ent.key(a.val), ent.key(b.val), a.val, b.val, bytes, 0);
}

__global__ void __launch_bounds__(256) tie_mark_kernel(SortBufs bufs, TieItem *ties, uint32_t tie_cap,
                                                       EntryArrays ent, const uint8_t *bytes, Ctl *ctl,
                                                       unsigned long long *rows_count, unsigned long long *rows_status,
                                                       uint32_t rows_status_words)
{
    const uint64_t n = ctl->n;
    {
        // what the rows kernel expects to be zero: its counts (at most one row per item), its
        // look-back words, its ticket
        const uint64_t gstride = (uint64_t)gridDim.x * blockDim.x, g0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (uint64_t i = g0; i < n; i += gstride) rows_count[i] = 0;
        for (uint64_t i = g0; i < rows_status_words; i += gstride) rows_status[i] = 0;
        if (g0 == 0) ctl->rows_ticket = 0;
    }
    SortItem *items = bufs.item[ctl->npasses & 1];
    const uint32_t lane = threadIdx.x & 31, lt = lanemask_lt();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t rounds = (n + stride - 1) / stride;
    uint32_t max_len = 0;
    for (uint64_t r = 0; r < rounds; ++r) {
        const uint64_t i = r * stride + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        bool big = false;
        uint32_t myval = 0;
        if (i < n) {
            const SortItem me = items[i];
            if ((me.lo & 0xFF) == 8) {
                // extent of my group, looking at most TIE_SMALL_GROUP items each way
                uint64_t lo_i = i, hi_i = i + 1;
                while (lo_i > 0 && i - lo_i < TIE_SMALL_GROUP) {
                    const SortItem o = items[lo_i - 1];
                    if (o.hi != me.hi || o.lo != me.lo) break;
                    --lo_i;
                }
                while (hi_i < n && hi_i - lo_i <= TIE_SMALL_GROUP) {
                    const SortItem o = items[hi_i];
                    if (o.hi != me.hi || o.lo != me.lo) break;
                    ++hi_i;
                }
                const uint32_t g = (uint32_t)(hi_i - lo_i);
                if (g > TIE_SMALL_GROUP) {
                    big = true;
                    myval = me.val;
                    max_len = max(max_len, key_len(ent.key(me.val)));
                } else if (g > 1 && lo_i == i) {
                    // head of a small group: insertion sort of the entry indices by full key
                    uint32_t v[TIE_SMALL_GROUP];
                    for (uint32_t a = 0; a < g; ++a) v[a] = items[i + a].val;
                    for (uint32_t a = 1; a < g; ++a) {
                        const uint32_t x = v[a];
                        const Key kx = ent.key(x);
                        uint32_t b = a;
                        while (b > 0 && key_less(kx, ent.key(v[b - 1]), x, v[b - 1], bytes, 1)) { v[b] = v[b - 1]; --b; }
                        v[b] = x;
                    }
                    for (uint32_t a = 0; a < g; ++a) items[i + a].val = v[a];
                    atomicAdd(&ctl->tie_small, 1u);
                }
            }
        }
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, big);
        if (!m) continue;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&ctl->n_tie, (unsigned int)__popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (big) {
            uint32_t k = base + __popc(m & lt);
            if (k < tie_cap) { ties[k].pos = (uint32_t)i; ties[k].val = myval; }
        }
    }
    // how many 7-byte rounds a full re-sort would need (run_lsd_ties): one atomic per warp
    for (int d = 16; d; d >>= 1) max_len = max(max_len, __shfl_xor_sync(0xFFFFFFFFu, max_len, d));
    if (lane == 0 && max_len) atomicMax(&ctl->max_tie_len, max_len);
}

// Many large tie groups (keys sharing long prefixes: bigrams, URLs, paths): instead of comparing
// whole keys pairwise, sort everything again least-significant round first -- sort word of round
// rmax, ..., of round 1, then (partition, round 0); every round is the same stable radix sort,
// so the result is the full order.  This kernel writes round `round`'s items into item[0] (from
// wherever the previous round left them) and collects the varying bits for the digit plan.
__global__ void sort_round_reset_kernel(Ctl *ctl)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) { ctl->or_lo = 0; ctl->and_lo = ~0ull; ctl->or_hi = 0; ctl->and_hi = ~0u; } // npasses stays: it says where the items are
    if (t < MAX_PASSES) ctl->tickets[t] = 0;
    for (uint32_t i = t; i < MAX_PASSES * RADIX_BINS; i += gridDim.x * blockDim.x) (&ctl->ghist[0][0])[i] = 0;
}

__global__ void __launch_bounds__(256) rekey_kernel(SortBufs bufs, uint32_t round, EntryArrays ent,
                                                    const uint8_t *bytes, unsigned long long max_keys, Ctl *ctl)
{
    const SortItem *src = bufs.item[ctl->npasses & 1];
    SortItem *dst = bufs.item[0];
    uint32_t hi_or = 0, hi_and = ~0u, pmin = ~0u, pmax = 0;
    uint64_t lo_or = 0, lo_and = ~0ull;
    bool any = false;
    const uint64_t n = ctl->n < max_keys ? ctl->n : max_keys;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        SortItem it = src[i];
        const ulonglong4 e = ld256(&ent.e[it.val]);
        Key key; key.k0 = e.x; key.k1 = e.y;
        it.lo = sort_chunk(key, bytes, round);
        it.hi = round == 0 ? (uint32_t)e.w : 0;
        dst[i] = it;
        hi_or |= it.hi; hi_and &= it.hi; lo_or |= it.lo; lo_and &= it.lo;
        any = true;
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any);
}

// Up to TIE_SMALL_MAX members of large groups: one CTA copies their keys into shared memory,
// orders the positions and the entries (full key compare) with bitonic networks and writes back.
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
            posv[i] = t.pos; sval[i] = t.val; spart[i] = ent.part(t.val); skey[i] = ent.key(t.val);
        }
    }
    __syncthreads();
    auto less = [&](uint32_t a, uint32_t b) -> bool { // a, b index the unsorted copies; >= nt is padding
        if (a >= nt || b >= nt) return a < nt && b >= nt;
        if (spart[a] != spart[b]) return spart[a] < spart[b];
        return key_less(skey[a], skey[b], sval[a], sval[b], bytes, 0);
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
    SortItem *items = bufs.item[ctl->npasses & 1];
    for (uint32_t i = threadIdx.x; i < nt; i += blockDim.x) items[posv[i]].val = sval[order[i]];
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

__global__ void __launch_bounds__(256) tie_writeback_kernel(SortBufs bufs, const TieItem *ties,
                                                            const uint32_t *posv, uint32_t nt, Ctl *ctl)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nt) bufs.item[ctl->npasses & 1][posv[i]].val = ties[i].val;
}

// ------------------------------------------------------------------ K6 + K7
struct RowArrays {
    uint32_t *part;
    unsigned long long *count;      // zeroed before rows_kernel
    unsigned long long *key_off;    // [n_out + 1]
    uint8_t *key_bytes;
    unsigned long long *part_begin; // [R + 1]
};

// look-back state of rows_kernel, one slot per tile
struct RowsStatus {
    unsigned long long *heads;      // flag << 62 | number of rows (aggregate or inclusive prefix)
    unsigned long long *bytes_agg;  // key bytes of the tile
    unsigned long long *bytes_inc;  // key bytes up to and including the tile
};

__device__ __forceinline__ void write_key_bytes(uint8_t *dst, const Key &key, uint32_t len, const uint8_t *bytes)
{
    if (!key_is_long(key)) {
        for (uint32_t i = 0; i < len && i < 8; ++i) dst[i] = (uint8_t)(key.k0 >> (8 * i));
        for (uint32_t i = 8; i < len; ++i) dst[i] = (uint8_t)(key.k1 >> (8 * (i - 8)));
    } else {
        const uint64_t off = key_off(key);
        for (uint32_t i = 0; i < len; i += 8) {
            uint64_t w = load8(bytes, off + i);
            uint32_t m = len - i < 8 ? len - i : 8;
            for (uint32_t j = 0; j < m; ++j) dst[i + j] = (uint8_t)(w >> (8 * j));
        }
    }
}

// Single pass over the sorted items: a position starts a row when its key differs from its
// left neighbour's; rows are numbered and their key bytes placed by two running sums carried
// across tiles with decoupled look-back; equal neighbours add their counts into the row
// (the segmented reduce; integer adds, so the order of the atomics does not matter).
__global__ void __launch_bounds__(ROWS_THREADS) rows_kernel(SortBufs bufs, EntryArrays ent, const uint8_t *bytes,
                                                           RowArrays rows, uint32_t num_reducers, uint32_t op,
                                                           unsigned long long key_bytes_cap, RowsStatus st, Ctl *ctl)
{
    __shared__ unsigned int wh[ROWS_THREADS / 32];
    __shared__ unsigned long long wb[ROWS_THREADS / 32];
    __shared__ unsigned long long s_prefix_h, s_prefix_b;
    __shared__ unsigned int s_tile;
    // tile id = atomic ticket: a tile waits only for tiles whose CTA has already started
    if (threadIdx.x == 0) s_tile = atomicAdd(&ctl->rows_ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t n = ctl->n;
    if (n == 0) {
        if (tile == 0) {
            for (uint32_t p = threadIdx.x; p <= num_reducers; p += ROWS_THREADS) rows.part_begin[p] = 0;
            if (threadIdx.x == 0) { ctl->n_out = 0; ctl->key_bytes = 0; rows.key_off[0] = 0; }
        }
        return;
    }
    if ((uint64_t)tile * ROWS_TILE >= n) return;
    const SortItem *items = bufs.item[ctl->npasses & 1];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t base = (uint64_t)tile * ROWS_TILE + (uint64_t)threadIdx.x * ROWS_ITEMS;

    SortItem it[ROWS_ITEMS];
    SortItem prev; prev.lo = 0; prev.hi = 0; prev.val = 0;
    if (base > 0 && base <= n) prev = items[base - 1];
#pragma unroll
    for (int j = 0; j < ROWS_ITEMS; ++j) {
        if (base + j < n) it[j] = items[base + j];
        else { it[j].lo = 0; it[j].hi = 0; it[j].val = 0; }
    }
    uint32_t headmask = 0, len[ROWS_ITEMS];
    Key key[ROWS_ITEMS];
    unsigned long long cnt[ROWS_ITEMS];
    uint32_t nh = 0;
    unsigned long long nb = 0;
#pragma unroll
    for (int j = 0; j < ROWS_ITEMS; ++j) {
        len[j] = 0; cnt[j] = 0;
        if (base + j >= n) continue;
        const SortItem &p = j == 0 ? prev : it[j - 1];
        bool head = base + j == 0 || p.hi != it[j].hi || p.lo != it[j].lo;
        const ulonglong4 e = ld256(&ent.e[it[j].val]); // key + count in one 32-byte gather
        key[j].k0 = e.x; key[j].k1 = e.y; cnt[j] = e.z;
        if (!head) head = !same_key(key[j], ent.key(p.val), bytes);
        if (head) { headmask |= 1u << j; len[j] = key_len(key[j]); ++nh; nb += len[j]; }
    }
    // block-wide exclusive sums of (rows, key bytes)
    uint32_t ih = nh;
    unsigned long long ib = nb;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t vh = __shfl_up_sync(0xFFFFFFFFu, ih, d);
        unsigned long long vb = __shfl_up_sync(0xFFFFFFFFu, ib, d);
        if (lane >= d) { ih += vh; ib += vb; }
    }
    if (lane == 31) { wh[warp] = ih; wb[warp] = ib; }
    __syncthreads();
    uint32_t oh = ih - nh, th = 0;
    unsigned long long ob = ib - nb, tb = 0;
#pragma unroll
    for (int w = 0; w < ROWS_THREADS / 32; ++w) {
        if (w < (int)warp) { oh += wh[w]; ob += wb[w]; }
        th += wh[w]; tb += wb[w];
    }
    // publish the tile aggregate, look back for the exclusive prefix (warp 0)
    constexpr unsigned long long F_AGG = 1ull << 62, F_INC = 2ull << 62, F_VAL = (1ull << 62) - 1;
    if (warp == 0) {
        volatile unsigned long long *vheads = st.heads, *vagg = st.bytes_agg, *vinc = st.bytes_inc;
        unsigned long long ph = 0, pb = 0;
        if (tile == 0) {
            if (lane == 0) { vinc[0] = tb; __threadfence(); vheads[0] = F_INC | th; }
        } else {
            if (lane == 0) { vagg[tile] = tb; __threadfence(); vheads[tile] = F_AGG | th; }
            int t = (int)tile - 1;
            for (uint32_t polls = 0;; ++polls) {
                if (polls > LOOKBACK_SPIN_LIMIT) { if (lane == 0) atomicOr(&ctl->err, ERR_LOOKBACK_TIMEOUT); break; }
                const int idx = t - (int)lane;
                unsigned long long w = idx >= 0 ? vheads[idx] : F_INC;
                const uint32_t flag = (uint32_t)(w >> 62);
                const uint32_t notready = __ballot_sync(0xFFFFFFFFu, flag == 0);
                const uint32_t incm = __ballot_sync(0xFFFFFFFFu, flag == 2);
                // lanes up to the first inclusive prefix take part
                const uint32_t stop = incm ? (uint32_t)__ffs((int)incm) - 1u : 31u;
                const uint32_t use = stop == 31 ? 0xFFFFFFFFu : ((2u << stop) - 1u);
                if (notready & use) continue; // a needed predecessor has not published yet
                __threadfence();
                unsigned long long b = 0, h = 0;
                if ((use >> lane) & 1u) {
                    h = w & F_VAL;
                    if (idx >= 0) b = flag == 2 ? vinc[idx] : vagg[idx];
                }
                for (int d = 16; d; d >>= 1) {
                    h += __shfl_xor_sync(0xFFFFFFFFu, h, d);
                    b += __shfl_xor_sync(0xFFFFFFFFu, b, d);
                }
                ph += h; pb += b;
                if (incm) break;
                t -= 32;
            }
            if (lane == 0) { vinc[tile] = pb + tb; __threadfence(); vheads[tile] = F_INC | (ph + th); }
        }
        if (lane == 0) { s_prefix_h = ph; s_prefix_b = pb; }
    }
    __syncthreads();
    unsigned long long o = s_prefix_h + oh, boff = s_prefix_b + ob;
    const unsigned long long tile_b0 = s_prefix_b;
    const bool overflow = tile_b0 + tb > key_bytes_cap;
    if (overflow && threadIdx.x == 0) atomicOr(&ctl->err, ERR_KEYS_FULL);
    // key bytes of the tile go through shared memory when they fit, so that global memory
    // sees 16-byte coalesced stores; the staging buffer keeps the destination's alignment
    __shared__ __align__(16) uint8_t sbytes[ROWS_STAGE + 16];
    const uint32_t mis = (uint32_t)(tile_b0 & 15);
    const bool staged = !overflow && tb <= ROWS_STAGE;
#pragma unroll
    for (int j = 0; j < ROWS_ITEMS; ++j) {
        const uint64_t pos = base + j;
        if (pos >= n) break;
        const SortItem &p = j == 0 ? prev : it[j - 1];
        const bool head = (headmask >> j) & 1u;
        if (head) {
            rows.part[o] = it[j].hi;
            rows.key_off[o] = boff;
            if (staged) write_key_bytes(sbytes + mis + (uint32_t)(boff - tile_b0), key[j], len[j], bytes);
            else if (!overflow) write_key_bytes(rows.key_bytes + boff, key[j], len[j], bytes);
            // partition starts: rows are partition-sorted
            const uint32_t p_first = pos == 0 ? 0u : p.hi + 1u;
            if (pos == 0 || p.hi != it[j].hi)
                for (uint32_t q = p_first; q <= it[j].hi && q <= num_reducers; ++q) rows.part_begin[q] = o;
            boff += len[j];
            ++o;
        }
        // a row whose neighbours on both sides are other rows needs no atomic
        const bool next_known = j + 1 < ROWS_ITEMS;
        const bool next_head = next_known && (pos + 1 >= n || ((headmask >> (j + 1)) & 1u));
        if (head && next_head) rows.count[o - 1] = cnt[j];
        else op_apply(&rows.count[o - 1], cnt[j], op);
        if (pos == n - 1) { // totals and sentinels
            ctl->n_out = o;
            ctl->key_bytes = boff;
            rows.key_off[o] = boff;
            for (uint32_t q = it[j].hi + 1u; q <= num_reducers; ++q) rows.part_begin[q] = o;
        }
    }
    if (staged) {
        __syncthreads();
        uint8_t *dst = rows.key_bytes + (tile_b0 - mis); // 16-byte aligned
        const uint32_t end = mis + (uint32_t)tb;
        for (uint32_t c = threadIdx.x * 16; c < end; c += ROWS_THREADS * 16) {
            if (c >= mis && c + 16 <= end) {
                *(uint4 *)(dst + c) = *(const uint4 *)(sbytes + c);
            } else { // first / last chunk shared with a neighbouring tile: bytes only
                for (uint32_t q = c < mis ? mis : c; q < c + 16 && q < end; ++q) dst[q] = sbytes[q];
            }
        }
    }
}

// rows of a MIN job hold complements until here (common.cuh, ReduceOp)
__global__ void __launch_bounds__(256) rows_decode_kernel(RowArrays rows, uint32_t op, const Ctl *ctl)
{
    const uint64_t n = ctl->n_out;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        rows.count[i] = op_decode(op, rows.count[i]);
}

} // namespace mr
