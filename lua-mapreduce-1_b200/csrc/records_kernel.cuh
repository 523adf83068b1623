// records_kernel.cuh -- map + combine for jobs whose map output is already (number key, number
// value) records: group-by-sum on uint32 keys (BASELINE config 5).
//
// In the reference such a job's mapfn emits (key, value) numbers, emit groups them per key
// (job.lua:83-97) and the combiner / reducer sums them (reducefn.lua:1-5).  Here one thread
// takes one record: heavy-hitter keys are summed in the CTA's shared-memory table (4-way
// buckets, 64-bit sums), everything else goes to the global table with the same 128-bit
// compare-and-swap insertion as the text path.  A key travels as a 4-byte inline key holding
// its big-endian bytes, so the byte-wise key order of the reduce side IS numeric order
// (Lua `<` on numbers, utils.lua:214) and nothing downstream needs to know about numbers.
#pragma once
#include "common.cuh"

namespace mr {

struct RecordsParams {
    const uint32_t *keys, *vals;
    uint64_t n;
    GEntry *gtab;
    uint32_t gmask;
    unsigned long long *records; // running record count
    unsigned int *err;
    uint32_t op;                 // ReduceOp
};

__device__ __forceinline__ Key key_of_u32(uint32_t k)
{
    Key key;
    key.k0 = __byte_perm(k, 0, 0x0123); // big-endian bytes: byte 0 = most significant
    key.k1 = 4ull << 56;
    return key;
}

__device__ __forceinline__ uint32_t u32_of_key(const Key &key) { return __byte_perm((uint32_t)key.k0, 0, 0x0123); }

constexpr unsigned long long REC_TAG = 1ull << 40; // marks an occupied shared-table slot (key 0 is valid)

template <int LOG2_SLOTS>
__global__ void __launch_bounds__(1024, 1) map_records_kernel(RecordsParams p)
{
    constexpr uint32_t SLOTS = 1u << LOG2_SLOTS;
    extern __shared__ __align__(16) uint8_t smem[];
    unsigned long long *tkeys = (unsigned long long *)smem;
    unsigned long long *tsum = tkeys + SLOTS;
    for (uint32_t i = threadIdx.x; i < SLOTS; i += blockDim.x) { tkeys[i] = 0; tsum[i] = 0; }
    __syncthreads();
    unsigned long long mine = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += stride) {
        const uint32_t k = __ldg(p.keys + i), v = __ldg(p.vals + i);
        ++mine;
        const unsigned long long sk = (unsigned long long)k | REC_TAG;
        const uint32_t bucket = (k * 0x9E3779B1u) >> (32 - (LOG2_SLOTS - 2));
        unsigned long long *bk = tkeys + bucket * 4;
        int way = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned long long cur = *(volatile unsigned long long *)&bk[q];
            if (cur == 0) cur = atomicCAS(&bk[q], 0ull, sk) == 0 ? sk : *(volatile unsigned long long *)&bk[q];
            if (cur == sk) { way = q; break; }
        }
        const unsigned long long x = op_encode(p.op, v);
        if (way >= 0) {
            unsigned long long *sum = &tsum[bucket * 4 + way];
            if (p.op == OP_SUM || p.op == OP_COUNT) {
                // a 64-bit shared-memory add is a compare-and-swap loop, and the heavy hitters send a
                // quarter of all records to four addresses: two native 32-bit adds instead (x < 2^32;
                // the low word's carry goes to the high word; the flush below reads them together)
                const unsigned int lo = (unsigned int)x;
                const unsigned int old = atomicAdd((unsigned int *)sum, lo);
                if (old + lo < old) atomicAdd((unsigned int *)sum + 1, 1u);
            } else {
                op_apply(sum, x, p.op);
            }
        } else {
            gtable_add(p.gtab, p.gmask, nullptr, key_of_u32(k), x, p.err, p.op);
        }
    }
    for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(p.records, mine);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < SLOTS; i += blockDim.x) {
        const unsigned long long sk = tkeys[i];
        if (sk) gtable_add(p.gtab, p.gmask, nullptr, key_of_u32((uint32_t)sk), tsum[i], p.err, p.op);
    }
}

// ------------------------------------------------------------------ (key, value) pairs with string keys
// What a reducer reads from the run files of other mappers (job.lua:253-265: every line of a run is
// `return <key>,{<values>}`, utils.lua:218-230): n keys as bytes + offsets and one value per key (the
// host has already reduced a line's value list to one value).  The pairs go into the same global table
// the tokeniser fills, so text splits and ingested runs of one job merge like runs of different mappers.
constexpr uint32_t ERR_EMPTY_KEY = 64u;      // joins the ERR_* bits (16 = ERR_PEER_TIMEOUT, 32 = ERR_LOOKBACK_TIMEOUT)

struct PairsParams {
    const uint8_t *arena;
    unsigned long long bytes_at, off_at, val_at, n;   // arena offsets of the key bytes, key offsets [n + 1], values [n]
    GEntry *gtab;
    uint32_t gmask, op, max_key;
    unsigned long long *pairs;
    unsigned int *err;
};

// hash of a long key of known length: what scan_long_token gives for the same bytes when they hold no
// whitespace (so that a key met in a text split and in a run file is the same table entry)
__host__ __device__ inline uint32_t hash_known_len(const uint8_t *bytes, uint64_t off, uint32_t len)
{
    uint64_t h = 0xCBF29CE484222325ull;
    for (uint32_t i = 0; i < len; i += 8) {
        const uint32_t nb = len - i < 8 ? len - i : 8;
        h = (h ^ (load8(bytes, off + i) & low_bytes_mask(nb))) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 32;
    }
    h ^= len;
    h *= 0xD6E8FEB86659FD93ull;
    return (uint32_t)(h >> 32);
}

__global__ void __launch_bounds__(256) map_pairs_kernel(PairsParams p)
{
    const unsigned long long *off = (const unsigned long long *)(p.arena + p.off_at);
    const unsigned long long *val = (const unsigned long long *)(p.arena + p.val_at);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long mine = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += stride) {
        const uint64_t ko = p.bytes_at + off[i];
        const uint64_t len64 = off[i + 1] - off[i];
        if (len64 == 0) { atomicOr(p.err, ERR_EMPTY_KEY); continue; }
        if (len64 > p.max_key) { atomicOr(p.err, ERR_TOKEN_TOO_LONG); continue; }
        const uint32_t len = (uint32_t)len64;
        Key key;
        if (len <= (uint32_t)INLINE_MAX) {
            key.k0 = load8(p.arena, ko) & low_bytes_mask(len < 8 ? len : 8);
            key.k1 = (len > 8 ? load8(p.arena, ko + 8) & low_bytes_mask(len - 8) : 0ull) | (unsigned long long)len << 56;
        } else {
            key = make_long_key(ko, len, hash_known_len(p.arena, ko, len));
        }
        gtable_add(p.gtab, p.gmask, p.arena, key, op_encode(p.op, val[i]), p.err, p.op);
        ++mine;
    }
    for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(p.pairs, mine);
}

} // namespace mr
