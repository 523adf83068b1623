// bigram_kernel.cuh -- map + combine for the bigram count (BASELINE config 4): the WordCount
// mapfn (mapfn.lua:3-9) with the key `w_i .. " " .. w_{i+1}` for neighbouring tokens of one line
// (io.lines cuts at '\n' only; every other %s byte separates tokens inside the line).
//
// One thread per token start: it finds its token's end, walks over the whitespace after it (a
// '\n' ends the line: no pair), finds the next token, and counts the pair.  Keys of up to 15 bytes
// are inline as everywhere else; longer keys are materialised in a device key pool the first time
// they are seen (bump allocation, bytes written before the 128-bit compare-and-swap publishes the
// key) and compared by hash, length and bytes.  There are ~10^7 distinct bigrams in a corpus
// whose word count has ~5*10^5 keys, so this path lives in the global table; the CTA-table
// staging of the word-count kernel is not used here.
#pragma once
#include "common.cuh"

namespace mr {

struct BigramParams {
    const uint8_t *arena;
    uint64_t begin, end;    // token starts in [begin, end) are this launch's first tokens
    uint64_t limit;         // bytes below this offset are defined (data or '\n' padding)
    GEntry *gtab;
    uint32_t gmask;
    uint32_t max_key;
    uint8_t *pool;
    unsigned long long *pool_used;
    unsigned long long pool_cap;
    unsigned long long *pairs;
    unsigned int *err;
};

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 0x20u || (c - 9u) < 5u; }

// byte k of the key `tok1 .. " " .. tok2`
__device__ __forceinline__ uint32_t bigram_byte(const uint8_t *a, uint64_t s1, uint32_t l1, uint64_t s2, uint32_t k)
{
    return k < l1 ? a[s1 + k] : (k == l1 ? 0x20u : a[s2 + (k - l1 - 1)]);
}

__global__ void __launch_bounds__(256) map_bigram_kernel(BigramParams p)
{
    const uint8_t *a = p.arena;
    unsigned long long mine = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = p.begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.end; i += stride) {
        if (is_ws(a[i])) continue;
        if (i > 0 && !is_ws(a[i - 1])) continue;          // not the first byte of its token
        uint64_t e1 = i + 1;
        while (e1 < p.limit && !is_ws(a[e1])) ++e1;
        uint64_t j = e1;
        bool pair = true;
        while (j < p.limit && is_ws(a[j])) { if (a[j] == 0x0Au) { pair = false; break; } ++j; }
        if (!pair || j >= p.limit) continue;               // end of line / of the staged data
        uint64_t e2 = j + 1;
        while (e2 < p.limit && !is_ws(a[e2])) ++e2;
        const uint64_t l1 = e1 - i, l2 = e2 - j, L = l1 + 1 + l2;
        ++mine;
        if (L > p.max_key) { atomicOr(p.err, ERR_TOKEN_TOO_LONG); continue; }
        Key key;
        if (L <= INLINE_MAX) {
            key.k0 = 0; key.k1 = (unsigned long long)L << 56;
            for (uint32_t k = 0; k < (uint32_t)L; ++k) {
                const unsigned long long c = bigram_byte(a, i, (uint32_t)l1, j, k);
                if (k < 8) key.k0 |= c << (8 * k); else key.k1 |= c << (8 * (k - 8));
            }
            gtable_add(p.gtab, p.gmask, nullptr, key, 1, p.err);
            continue;
        }
        // long key: hash of the logical bytes, then probe; the key's bytes go to the pool on first sight
        unsigned long long h64 = 0xCBF29CE484222325ull;
        for (uint32_t k = 0; k < (uint32_t)L; ++k) h64 = (h64 ^ bigram_byte(a, i, (uint32_t)l1, j, k)) * 0x100000001B3ull;
        const uint32_t h = (uint32_t)(h64 ^ (h64 >> 32));
        const unsigned long long k1 = TAG_LONG | h;
        uint32_t slot = mix32(h * 0x9E3779B1u + 0x7F4A7C15u) & p.gmask; // = hash_key() of a long key
        unsigned long long my_off = ~0ull;
        bool done = false;
        for (uint32_t probes = 0; probes < GT_MAX_PROBES && !done; ++probes) {
            ulonglong2 v = __ldcg((const ulonglong2 *)&p.gtab[slot]);
            if (v.y == 0) {
                if (my_off == ~0ull) {
                    my_off = atomicAdd(p.pool_used, (unsigned long long)L);
                    if (my_off + L > p.pool_cap) { atomicOr(p.err, ERR_KEYS_FULL); done = true; break; }
                    for (uint32_t k = 0; k < (uint32_t)L; ++k) p.pool[my_off + k] = (uint8_t)bigram_byte(a, i, (uint32_t)l1, j, k);
                    __threadfence();                        // bytes before the key becomes visible
                }
                Key mykey;
                mykey.k0 = (my_off << LONG_LEN_BITS) | L;
                mykey.k1 = k1;
                const Key old = cas128(&p.gtab[slot], mykey);
                if (old.k1 == 0) { atomicAdd(&p.gtab[slot].cnt, 1ull); done = true; break; }
                v.x = old.k0; v.y = old.k1;
            }
            if (v.y == k1 && (v.x & LONG_LEN_MASK) == L) {
                const volatile uint8_t *q = p.pool + (v.x >> LONG_LEN_BITS);
                bool same = true;
                for (uint32_t k = 0; k < (uint32_t)L && same; ++k) same = q[k] == bigram_byte(a, i, (uint32_t)l1, j, k);
                if (same) { atomicAdd(&p.gtab[slot].cnt, 1ull); done = true; break; }
            }
            slot = (slot + 1) & p.gmask;
        }
        if (!done) atomicOr(p.err, ERR_TABLE_FULL);
    }
    for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(p.pairs, mine);
}

} // namespace mr
