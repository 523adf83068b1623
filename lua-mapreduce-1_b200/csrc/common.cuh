// common.cuh -- shared device helpers: key words, whitespace SWAR, hashing, the
// reference partitioner in binary64, global hash-table insertion.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mr {

// ---------------------------------------------------------------- key words
// Every distinct key is identified by one 64-bit word.
//   short key (1..7 bytes): bytes little-endian in bits 0..55, length in bits 56..62.
//                           The word IS the key, so equality is exact.
//   long key  (>= 8 bytes): bit 63 | byte offset (37 bits) << 26 | length (26 bits);
//                           the offset points at one occurrence in the byte space
//                           (input arena on the map side, receive buffer after a
//                           shuffle); equality compares the bytes.
// 0 is never a valid word (length >= 1), so it marks an empty table slot.
constexpr uint64_t KEY_LONG = 1ull << 63;
constexpr int LONG_LEN_BITS = 26;
constexpr uint64_t LONG_LEN_MASK = (1ull << LONG_LEN_BITS) - 1;
constexpr uint64_t MAX_BYTE_SPACE = 1ull << 37;

__host__ __device__ __forceinline__ bool key_is_long(uint64_t k) { return (k >> 63) != 0; }
__host__ __device__ __forceinline__ uint32_t key_len(uint64_t k)
{
    return key_is_long(k) ? (uint32_t)(k & LONG_LEN_MASK) : (uint32_t)(k >> 56);
}
__host__ __device__ __forceinline__ uint64_t key_off(uint64_t k)
{
    return (k & ~KEY_LONG) >> LONG_LEN_BITS;
}
__host__ __device__ __forceinline__ uint64_t make_long_key(uint64_t off, uint32_t len)
{
    return KEY_LONG | (off << LONG_LEN_BITS) | len;
}

struct GEntry {
    unsigned long long key;
    unsigned long long cnt;
};

// error bits accumulated on the device (Ctl::err)
enum : uint32_t {
    ERR_TABLE_FULL = 1u,
    ERR_TOKEN_TOO_LONG = 2u,
    ERR_KEYS_FULL = 4u,
};

// ---------------------------------------------------------------- whitespace
// C-locale isspace (Lua's %s, mapfn.lua:5): 0x20 and 0x09..0x0D.  Returns 0x80 in
// every byte lane of x that is whitespace; exact for every lane (no borrow tricks).
__device__ __forceinline__ uint32_t space_flags(uint32_t x)
{
    uint32_t y = x & 0x7F7F7F7Fu;
    uint32_t eq20 = ~((y ^ 0x20202020u) + 0x7F7F7F7Fu); // bit7 set iff low7 == 0x20
    uint32_t ge9 = y + 0x77777777u;                      // bit7 set iff low7 >= 9
    uint32_t lt14 = ~(y + 0x72727272u);                  // bit7 set iff low7 < 14
    return (eq20 | (ge9 & lt14)) & ~x & 0x80808080u;
}

// 4-bit mask (bit i = byte i of x is NOT whitespace)
__device__ __forceinline__ uint32_t nonspace_nibble(uint32_t x)
{
    uint32_t f = (~space_flags(x)) & 0x80808080u;
    return (((f >> 7) * 0x01020408u) >> 24) & 0xFu;
}

// ---------------------------------------------------------------- hashing
__device__ __forceinline__ uint32_t hash_word(uint64_t key)
{
    uint32_t lo = (uint32_t)key, hi = (uint32_t)(key >> 32);
    uint32_t h = (lo * 0x9E3779B1u) ^ (hi * 0x85EBCA77u);
    h ^= h >> 15; h *= 0x2C1B3C6Du;
    h ^= h >> 12; h *= 0x297A2D39u;
    h ^= h >> 15;
    return h;
}

// 8 bytes at an arbitrary byte offset, little-endian.  May read up to 15 bytes past
// off; every byte space carries that much slack.
__device__ __forceinline__ uint64_t load8(const uint8_t *base, uint64_t off)
{
    const unsigned long long *p = (const unsigned long long *)(base + (off & ~7ull));
    uint32_t sh = (uint32_t)(off & 7) * 8;
    uint64_t w0 = __ldg(p);
    if (sh == 0) return w0;
    uint64_t w1 = __ldg(p + 1);
    return (w0 >> sh) | (w1 << (64 - sh));
}

__device__ __forceinline__ uint64_t low_bytes_mask(uint32_t nbytes) // nbytes in 0..8
{
    return nbytes >= 8 ? ~0ull : ((1ull << (8 * nbytes)) - 1);
}

__device__ __forceinline__ uint64_t bswap64(uint64_t v)
{
    uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
    return ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
}

__device__ inline bool bytes_equal(const uint8_t *base, uint64_t a, uint64_t b, uint32_t len)
{
    if (a == b) return true;
    for (uint32_t i = 0; i < len; i += 8) {
        uint64_t m = low_bytes_mask(len - i);
        if ((load8(base, a + i) ^ load8(base, b + i)) & m) return false;
    }
    return true;
}

// unsigned-byte lexicographic compare of two long keys, shorter first on equal prefix
// (Lua string `<`, utils.lua:126 / utils.lua:214)
__device__ inline int long_key_compare(const uint8_t *base, uint64_t ka, uint64_t kb)
{
    uint32_t la = key_len(ka), lb = key_len(kb);
    uint64_t oa = key_off(ka), ob = key_off(kb);
    uint32_t n = la < lb ? la : lb;
    for (uint32_t i = 0; i < n; i += 8) {
        uint64_t m = low_bytes_mask(n - i);
        uint64_t wa = bswap64(load8(base, oa + i) & m), wb = bswap64(load8(base, ob + i) & m);
        if (wa != wb) return wa < wb ? -1 : 1;
    }
    return la < lb ? -1 : (la > lb ? 1 : 0);
}

// ---------------------------------------------------------------- partitioner
// examples/WordCount/partitionfn.lua:10-15 in IEEE binary64 with Lua 5.2's
// `a % b = a - floor(a/b)*b`.  h * prime exceeds 2^53 and rounds (nearest even);
// dividing and multiplying by 2^32 are exact, so one rounding per byte, as in Lua.
struct FnvLua52 {
    double h;
    __device__ __forceinline__ FnvLua52() : h(2166136261.0) {}
    __device__ __forceinline__ void byte(uint32_t c)
    {
        double p = __dmul_rn(h, 16777619.0);
        double q = floor(__dmul_rn(p, 1.0 / 4294967296.0));
        double m = __fma_rn(-q, 4294967296.0, p);
        uint32_t u = __double2uint_rz(m);
        h = (double)(u ^ c);
    }
    __device__ __forceinline__ uint32_t partition(uint32_t num_reducers) const
    {
        double r = (double)num_reducers;
        double q = floor(__ddiv_rn(h, r));
        return __double2uint_rz(__fma_rn(-q, r, h));
    }
};

__device__ inline uint32_t partition_of_key(uint64_t key, const uint8_t *bytes, uint32_t num_reducers)
{
    FnvLua52 f;
    if (!key_is_long(key)) {
        uint32_t len = (uint32_t)(key >> 56);
        for (uint32_t i = 0; i < len; ++i) f.byte((uint32_t)(key >> (8 * i)) & 0xFFu);
    } else {
        uint32_t len = key_len(key);
        uint64_t off = key_off(key);
        for (uint32_t i = 0; i < len; i += 8) {
            uint64_t w = load8(bytes, off + i);
            uint32_t n = len - i < 8 ? len - i : 8;
            for (uint32_t j = 0; j < n; ++j) f.byte((uint32_t)(w >> (8 * j)) & 0xFFu);
        }
    }
    return f.partition(num_reducers);
}

// Sort word of a key for ordering round r (7 key bytes per round, big-endian, then
// min(remaining length, 8)): numeric order of the words of round 0, then round 1,
// ... is unsigned-byte lexicographic order with the shorter key first.
__device__ inline uint64_t sort_chunk(uint64_t key, const uint8_t *bytes, uint32_t round)
{
    if (!key_is_long(key)) {
        if (round) return 0;
        return bswap64(key); // byte0 on top ... byte6, then the length byte (<= 7)
    }
    uint32_t len = key_len(key);
    uint64_t skip = 7ull * round;
    if (skip >= len) return 0;
    uint32_t rem = len - (uint32_t)skip;
    uint64_t w = load8(bytes, key_off(key) + skip) & low_bytes_mask(rem < 7 ? rem : 7);
    return (bswap64(w) & ~0xFFull) | (rem < 8 ? rem : 8);
}

// ---------------------------------------------------------------- global table
// Open addressing, linear probing, 16-byte slots {key word, count}.  Keys are
// claimed with a 64-bit CAS and never change afterwards; counts are added with
// fire-and-forget 64-bit reductions.
constexpr uint32_t GT_MAX_PROBES = 1u << 14;

__device__ inline void gtable_add_short(GEntry *gt, uint32_t gmask, uint64_t key, uint64_t n,
                                        unsigned int *err)
{
    uint32_t slot = hash_word(key) & gmask;
    for (uint32_t probes = 0; probes < GT_MAX_PROBES; ++probes) {
        unsigned long long k = __ldcg(&gt[slot].key);
        if (k == 0) {
            k = atomicCAS(&gt[slot].key, 0ull, (unsigned long long)key);
            if (k == 0) k = key;
        }
        if (k == key) {
            atomicAdd(&gt[slot].cnt, (unsigned long long)n);
            return;
        }
        slot = (slot + 1) & gmask;
    }
    atomicOr(err, ERR_TABLE_FULL);
}

// key = long reference into `bytes`; h = hash of its bytes
__device__ inline void gtable_add_long(GEntry *gt, uint32_t gmask, const uint8_t *bytes,
                                       uint64_t key, uint32_t h, uint64_t n, unsigned int *err)
{
    uint32_t slot = h & gmask;
    uint32_t len = key_len(key);
    for (uint32_t probes = 0; probes < GT_MAX_PROBES; ++probes) {
        unsigned long long k = __ldcg(&gt[slot].key);
        if (k == 0) {
            k = atomicCAS(&gt[slot].key, 0ull, (unsigned long long)key);
            if (k == 0) k = key;
        }
        if (key_is_long(k) && key_len(k) == len &&
            bytes_equal(bytes, key_off(k), key_off(key), len)) {
            atomicAdd(&gt[slot].cnt, (unsigned long long)n);
            return;
        }
        slot = (slot + 1) & gmask;
    }
    atomicOr(err, ERR_TABLE_FULL);
}

// Length (up to the next whitespace byte) and hash of the token starting at off.
// Returns 0 if the token is longer than max_len.
__device__ inline uint32_t scan_long_token(const uint8_t *bytes, uint64_t off, uint32_t max_len,
                                           uint32_t *hash_out)
{
    uint64_t h = 0xCBF29CE484222325ull;
    uint32_t len = 0;
    for (;;) {
        uint64_t w = load8(bytes, off + len);
        uint32_t f0 = space_flags((uint32_t)w), f1 = space_flags((uint32_t)(w >> 32));
        uint32_t nb = 8;
        if (f0) nb = (uint32_t)(__ffs((int)f0) - 8) >> 3;
        else if (f1) nb = 4 + ((uint32_t)(__ffs((int)f1) - 8) >> 3);
        if (nb) {
            h = (h ^ (w & low_bytes_mask(nb))) * 0x9E3779B97F4A7C15ull;
            h ^= h >> 32;
        }
        len += nb;
        if (nb < 8) break;
        if (len > max_len) return 0;
    }
    if (len > max_len) return 0;
    h ^= len;
    h *= 0xD6E8FEB86659FD93ull;
    *hash_out = (uint32_t)(h >> 32);
    return len;
}

__device__ __forceinline__ uint32_t lanemask_lt()
{
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

} // namespace mr
