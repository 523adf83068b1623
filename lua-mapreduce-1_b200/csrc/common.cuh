// common.cuh -- shared device helpers: key words, whitespace SWAR, hashing, the
// reference partitioner in binary64, global hash-table insertion.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mr {

// ---------------------------------------------------------------- keys
// Every distinct key is identified by 128 bits (k0, k1).
//   inline key (1..15 bytes): k0 = bytes 0..7, k1 bits 0..55 = bytes 8..14 (little-endian,
//                             zero padded), k1 bits 56..63 = length.  The pair IS the key,
//                             so equality is exact and no input byte is ever re-read.
//   long key (>= 16 bytes):   k1 = 0xFF << 56 | 32-bit hash of the bytes,
//                             k0 = byte offset << 26 | length (26 bits); the offset points
//                             at one occurrence in the byte space (input arena on the map
//                             side, receive buffer after a shuffle); equality = same hash,
//                             same length, same bytes.
// k1 == 0 never occurs for a valid key (length >= 1), so it marks an empty table slot.
constexpr int INLINE_MAX = 15;
constexpr uint64_t TAG_LONG = 0xFFull << 56;
constexpr uint64_t LOW56 = (1ull << 56) - 1;
constexpr int LONG_LEN_BITS = 26;
constexpr uint64_t LONG_LEN_MASK = (1ull << LONG_LEN_BITS) - 1;
constexpr uint64_t MAX_BYTE_SPACE = 1ull << 37;

struct Key {
    unsigned long long k0, k1;
};

__host__ __device__ __forceinline__ bool key_is_long(const Key &k) { return (k.k1 >> 56) == 0xFF; }
__host__ __device__ __forceinline__ uint32_t key_len(const Key &k)
{
    return key_is_long(k) ? (uint32_t)(k.k0 & LONG_LEN_MASK) : (uint32_t)(k.k1 >> 56);
}
__host__ __device__ __forceinline__ uint64_t key_off(const Key &k) { return k.k0 >> LONG_LEN_BITS; }
__host__ __device__ __forceinline__ Key make_long_key(uint64_t off, uint32_t len, uint32_t hash)
{
    Key k;
    k.k0 = (off << LONG_LEN_BITS) | len;
    k.k1 = TAG_LONG | hash;
    return k;
}

struct __align__(32) GEntry {
    unsigned long long k0, k1;
    unsigned long long cnt;
    unsigned long long pad;
};

// error bits accumulated on the device (Ctl::err)
enum : uint32_t {
    ERR_TABLE_FULL = 1u,
    ERR_TOKEN_TOO_LONG = 2u,
    ERR_KEYS_FULL = 4u,
    ERR_SHUFFLE_FULL = 8u,
};

// ---------------------------------------------------------------- both sides of the fence
// The partitioner, the sort words and the owner map are also compiled for the host
// (tests/host_helpers.cu), so that the very code the kernels run is unit-tested on a machine
// without a GPU.  The intrinsics they use, spelled for both sides; in device code each
// is exactly the intrinsic it names.
__host__ __device__ __forceinline__ unsigned long long ld_ro64(const unsigned long long *p)
{
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
__host__ __device__ __forceinline__ uint32_t bswap32(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return __byte_perm(x, 0, 0x0123);
#else
    return __builtin_bswap32(x);
#endif
}
__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b)
{
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (uint32_t)(((unsigned long long)a * b) >> 32);
#endif
}
__host__ __device__ __forceinline__ int ffs64(unsigned long long x) // 1-based index of the lowest set bit, 0 if none
{
#ifdef __CUDA_ARCH__
    return __ffsll((long long)x);
#else
    return __builtin_ffsll((long long)x);
#endif
}
__host__ __device__ __forceinline__ int ffs32(uint32_t x) // 1-based index of the lowest set bit, 0 if none
{
#ifdef __CUDA_ARCH__
    return __ffs((int)x);
#else
    return __builtin_ffs((int)x);
#endif
}
__host__ __device__ __forceinline__ int clz32(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// ---------------------------------------------------------------- whitespace
// C-locale isspace (Lua's %s, mapfn.lua:5): 0x20 and 0x09..0x0D.  Returns 0x80 in
// every byte lane of x that is whitespace; exact for every lane (no borrow tricks).
__host__ __device__ __forceinline__ uint32_t space_flags(uint32_t x)
{
    uint32_t y = x & 0x7F7F7F7Fu;
    uint32_t eq20 = ~((y ^ 0x20202020u) + 0x7F7F7F7Fu); // bit7 set iff low7 == 0x20
    uint32_t ge9 = y + 0x77777777u;                      // bit7 set iff low7 >= 9
    uint32_t lt14 = ~(y + 0x72727272u);                  // bit7 set iff low7 < 14
    return (eq20 | (ge9 & lt14)) & ~x & 0x80808080u;
}

// 4-bit mask (bit i = byte i of x is NOT whitespace)
__host__ __device__ __forceinline__ uint32_t nonspace_nibble(uint32_t x)
{
    uint32_t f = (~space_flags(x)) & 0x80808080u;
    return (((f >> 7) * 0x01020408u) >> 24) & 0xFu;
}

// ---------------------------------------------------------------- token bounds from masks
// m: bit i = byte i of a lane's 32 bytes is a token byte; nxt: the same for the 32 bytes that
// follow; carry: 1 if the byte before byte 0 is a token byte.  A token starts where a token byte
// follows a non-token byte; its class comes from mask algebra alone: bit i of t8 says bytes
// i..i+7 are all token bytes (length >= 8), t16 the same for 16.
struct TokenMasks {
    unsigned long long w;   // m | nxt << 32
    uint32_t starts;        // token starts among the lane's 32 bytes
    uint32_t s_long;        // ... of tokens of 16+ bytes
    uint32_t s_med;         // ... of 8..15 bytes
    uint32_t s_short;       // ... of 1..7 bytes
};
__host__ __device__ __forceinline__ TokenMasks token_masks(uint32_t m, uint32_t nxt, uint32_t carry)
{
    TokenMasks r;
    r.w = (unsigned long long)m | ((unsigned long long)nxt << 32);
    r.starts = m & ~((m << 1) | carry);
    unsigned long long t = r.w & (r.w >> 1);
    t &= t >> 2;
    const unsigned long long t8 = t & (t >> 4);
    const unsigned long long t16 = t8 & (t8 >> 8);
    r.s_long = r.starts & (uint32_t)t16;
    r.s_med = r.starts & (uint32_t)t8 & ~(uint32_t)t16;
    r.s_short = r.starts & ~(uint32_t)t8;
    return r;
}
// length (1..15) of the short or medium token that starts at bit `bit` of w's low word
__host__ __device__ __forceinline__ uint32_t token_len(unsigned long long w, uint32_t bit)
{
#ifdef __CUDA_ARCH__
    return __ffs((int)~(uint32_t)(w >> bit)) - 1;
#else
    return (uint32_t)__builtin_ffs((int)~(uint32_t)(w >> bit)) - 1;
#endif
}

// ---------------------------------------------------------------- inline keys from staged words
// x with the bytes at index >= n cleared; n may be <= 0 (clears all) or >= 4 (keeps all).
// PTX shl clamps shift amounts above 31, which C does not.
__host__ __device__ __forceinline__ uint32_t keep_low_bytes(uint32_t x, int n)
{
#ifdef __CUDA_ARCH__
    uint32_t hi;
    asm("shl.b32 %0, %1, %2;" : "=r"(hi) : "r"(0xFFFFFFFFu), "r"((uint32_t)max(n, 0) * 8u));
    return x & ~hi;
#else
    return n <= 0 ? 0u : n >= 4 ? x : x & ((1u << (8 * n)) - 1u);
#endif
}
// the 32 bits that start `sh` bits (0, 8, 16, 24) into the 64-bit pair (lo, hi)
__host__ __device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)
{
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, sh);
#else
    sh &= 31;
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
// a token of 1..7 bytes that starts bs/8 bytes into the aligned words x0, x1, x2:
// w0 = bytes 0..3, w1 = bytes 4..6 | length << 24 (the two halves of an inline key's words)
struct ShortKey { uint32_t w0, w1; };
__host__ __device__ __forceinline__ ShortKey short_key(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t bs, uint32_t len)
{
    ShortKey k;
    k.w0 = keep_low_bytes(funnel_r(x0, x1, bs), (int)len);
    k.w1 = keep_low_bytes(funnel_r(x1, x2, bs), (int)len - 4) | (len << 24);
    return k;
}
// a token of 8..15 bytes from five aligned words: the 128-bit inline key
__host__ __device__ __forceinline__ Key medium_key(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t x4,
                                                   uint32_t bs, uint32_t len)
{
    Key key;
    key.k0 = (uint64_t)funnel_r(x0, x1, bs) | ((uint64_t)funnel_r(x1, x2, bs) << 32);
    const uint32_t w2 = keep_low_bytes(funnel_r(x2, x3, bs), (int)len - 8);
    const uint32_t w3 = keep_low_bytes(funnel_r(x3, x4, bs), (int)len - 12) | (len << 24);
    key.k1 = (uint64_t)w2 | ((uint64_t)w3 << 32);
    return key;
}

// ---------------------------------------------------------------- hashing
__device__ __forceinline__ uint32_t mix32(uint32_t h)
{
    h ^= h >> 15; h *= 0x2C1B3C6Du;
    h ^= h >> 12; h *= 0x297A2D39u;
    h ^= h >> 15;
    return h;
}

// hash of a short key held in one word (shared-memory table)
__device__ __forceinline__ uint32_t hash_word(uint64_t w)
{
    return mix32(((uint32_t)w * 0x9E3779B1u) ^ ((uint32_t)(w >> 32) * 0x85EBCA77u));
}

// hash of a full key (global table).  A long key hashes by its byte hash only, so that
// every occurrence of the same bytes lands on the same probe sequence.
__device__ __forceinline__ uint32_t hash_key(const Key &k)
{
    if (key_is_long(k)) return mix32((uint32_t)k.k1 * 0x9E3779B1u + 0x7F4A7C15u);
    uint32_t h = ((uint32_t)k.k0 * 0x9E3779B1u) ^ ((uint32_t)(k.k0 >> 32) * 0x85EBCA77u);
    h ^= ((uint32_t)k.k1 * 0xC2B2AE3Du) ^ ((uint32_t)(k.k1 >> 32) * 0x27D4EB2Fu);
    return mix32(h);
}

// 8 bytes at an arbitrary byte offset, little-endian.  May read up to 15 bytes past
// off; every byte space carries that much slack.
__host__ __device__ __forceinline__ uint64_t load8(const uint8_t *base, uint64_t off)
{
    const unsigned long long *p = (const unsigned long long *)(base + (off & ~7ull));
    uint32_t sh = (uint32_t)(off & 7) * 8;
    uint64_t w0 = ld_ro64(p);
    if (sh == 0) return w0;
    uint64_t w1 = ld_ro64(p + 1);
    return (w0 >> sh) | (w1 << (64 - sh));
}

__host__ __device__ __forceinline__ uint64_t low_bytes_mask(uint32_t nbytes) // nbytes in 0..8
{
    return nbytes >= 8 ? ~0ull : ((1ull << (8 * nbytes)) - 1);
}

__host__ __device__ __forceinline__ uint64_t bswap64(uint64_t v)
{
    uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
    return ((uint64_t)bswap32(lo) << 32) | bswap32(hi);
}

__host__ __device__ inline bool bytes_equal(const uint8_t *base, uint64_t a, uint64_t b, uint32_t len)
{
    if (a == b) return true;
    for (uint32_t i = 0; i < len; i += 8) {
        uint64_t m = low_bytes_mask(len - i);
        if ((load8(base, a + i) ^ load8(base, b + i)) & m) return false;
    }
    return true;
}

__host__ __device__ __forceinline__ bool same_key(const Key &a, const Key &b, const uint8_t *bytes)
{
    if (a.k1 != b.k1) return false;       // length / hash differ
    if (a.k0 == b.k0) return true;
    if (!key_is_long(a)) return false;
    uint32_t la = key_len(a);
    return la == key_len(b) && bytes_equal(bytes, key_off(a), key_off(b), la);
}

// Sort word of a key for ordering round r: key bytes 7r .. 7r+6 big-endian in the top 56
// bits, min(remaining length, 8) in the low byte.  Numeric order of the words of round 0,
// then round 1, ... is unsigned-byte lexicographic order with the shorter key first
// (Lua string `<`, utils.lua:126 / utils.lua:214).
__host__ __device__ inline uint64_t sort_chunk(const Key &k, const uint8_t *bytes, uint32_t round)
{
    const uint32_t len = key_len(k);
    const uint32_t skip = 7u * round;
    if (skip >= len) return 0;
    const uint32_t rem = len - skip;
    uint64_t w;
    if (!key_is_long(k)) {
        const uint64_t hi = k.k1 & LOW56;
        if (round == 0) w = k.k0;
        else if (round == 1) w = (k.k0 >> 56) | (hi << 8);
        else w = hi >> 48;               // byte 14
    } else {
        w = load8(bytes, key_off(k) + skip);
    }
    w &= low_bytes_mask(rem < 7 ? rem : 7);
    return (bswap64(w) & ~0xFFull) | (rem < 8 ? rem : 8);
}

// ---------------------------------------------------------------- owner rank of a key
// Which rank reduces a key.  Without splitters: partition mod world (every result.P<p> lives on
// one rank, as one reduce job owns one partition in the reference, server.lua:316-324).  With
// world-1 splitters (part, sort word of round 0), ascending: rank = number of splitters <= the
// key's (part, word) -- contiguous ranges of the global order "partition, then key", so that the
// ranks hold equal shares whatever the partition count, each result.P<p> being the concatenation
// of at most a few ranks' slices in rank order.  Keys equal in (part, word) stay together.
struct OwnerMap {
    const unsigned long long *word; // [n] splitter sort words
    const uint32_t *part;           // [n] splitter partitions
    uint32_t n;                     // 0 or world - 1
    uint32_t world;
};

__host__ __device__ __forceinline__ uint32_t owner_of(const OwnerMap &m, uint32_t part, uint64_t w /* sort_chunk(key, bytes, 0) */)
{
    if (m.n == 0) return part % m.world;
    uint32_t lo = 0, hi = m.n; // first splitter greater than (part, w)
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        const uint32_t sp = m.part[mid];
        const bool le = sp < part || (sp == part && m.word[mid] <= w);
        if (le) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ---------------------------------------------------------------- partitioner
// examples/WordCount/partitionfn.lua:10-15.  Lua 5.2 numbers are IEEE binary64 and
// `a % b = a - floor(a/b)*b`: h * prime can reach 2^57, so the product is rounded to 53
// significant bits (nearest, ties to even) before `% 2^32` keeps its low 32 bits -- this is
// NOT uint32 FNV-1.  The rounding is done here with integers (bit-for-bit the same result
// as the double arithmetic: oracle/mr_oracle.c checks both forms against each other, and
// the host-side mr_partition_of uses real doubles).  `h % NUM_REDUCERS` on integer-valued
// doubles below 2^32 equals the integer remainder.
struct FnvLua52 {
    uint32_t h;
    __host__ __device__ __forceinline__ FnvLua52() : h(2166136261u) {}
    __host__ __device__ __forceinline__ void byte(uint32_t c)
    {
        // Only the low 32 bits of the rounded product survive `% 2^32`, and the rounding touches
        // at most the low 4 bits plus a carry that runs upwards, so it is done on the low word;
        // the high word only says how many bits are dropped: p < 2^57, top bit = 63 - clz(p),
        // s = top - 52 = 11 - clz(high word) when that is positive.
        uint32_t lo = h * 16777619u;
        const uint32_t hi = mulhi32(h, 16777619u);
        const int s = 11 - clz32(hi);
        if (s > 0) {
            // round to nearest, ties to even: adding half-1+odd(quotient) carries exactly when it must
            const uint32_t add = (1u << (s - 1)) - 1u + ((lo >> s) & 1u);
            lo = ((lo + add) >> s) << s;
        }
        h = lo ^ c;
    }
    __host__ __device__ __forceinline__ uint32_t partition(uint32_t num_reducers) const { return h % num_reducers; }
};

__host__ __device__ inline uint32_t partition_of_key(const Key &k, const uint8_t *bytes, uint32_t num_reducers)
{
    FnvLua52 f;
    const uint32_t len = key_len(k);
    if (!key_is_long(k)) {
        const uint32_t n0 = len < 8 ? len : 8;
        for (uint32_t i = 0; i < n0; ++i) f.byte((uint32_t)(k.k0 >> (8 * i)) & 0xFFu);
        for (uint32_t i = 8; i < len; ++i) f.byte((uint32_t)(k.k1 >> (8 * (i - 8))) & 0xFFu);
    } else {
        const uint64_t off = key_off(k);
        for (uint32_t i = 0; i < len; i += 8) {
            uint64_t w = load8(bytes, off + i);
            uint32_t n = len - i < 8 ? len - i : 8;
            for (uint32_t j = 0; j < n; ++j) f.byte((uint32_t)(w >> (8 * j)) & 0xFFu);
        }
    }
    return f.partition(num_reducers);
}

// ---------------------------------------------------------------- reducers
// What is done with the values of one key (the task module's reducefn/combinerfn; all of them are
// associative, commutative and idempotent in the sense of job.lua:104-106).  Tables are cleared to
// zero, so every running value is kept in a form whose identity is 0: a sum as it is, a maximum as it
// is, a minimum as its bitwise complement (then "min" is a max of complements).
enum ReduceOp : uint32_t { OP_SUM = 0, OP_MIN = 1, OP_MAX = 2, OP_COUNT = 3 };

__host__ __device__ __forceinline__ unsigned long long op_encode(uint32_t op, unsigned long long v)
{
    return op == OP_COUNT ? 1ull : (op == OP_MIN ? ~v : v);
}
__host__ __device__ __forceinline__ unsigned long long op_decode(uint32_t op, unsigned long long x)
{
    return op == OP_MIN ? ~x : x;
}
// merge an encoded value into a running encoded value
__device__ __forceinline__ void op_apply(unsigned long long *addr, unsigned long long x, uint32_t op)
{
    if (op == OP_SUM || op == OP_COUNT) atomicAdd(addr, x);
    else atomicMax(addr, x);
}

// ---------------------------------------------------------------- global table
// Open addressing, linear probing, 32-byte slots {k0, k1, count, -}.  A key is claimed
// with ONE 128-bit compare-and-swap on (k0, k1) and never changes afterwards; counts
// are added with fire-and-forget 64-bit reductions.
constexpr uint32_t GT_MAX_PROBES = 1u << 14;

__device__ __forceinline__ Key cas128(GEntry *slot, const Key &desired)
{
    Key old;
    asm volatile(
        "{\n\t"
        ".reg .b128 cmp, des, res;\n\t"
        "mov.b128 cmp, {%2, %3};\n\t"
        "mov.b128 des, {%4, %5};\n\t"
        "atom.global.cas.b128 res, [%6], cmp, des;\n\t"
        "mov.b128 {%0, %1}, res;\n\t"
        "}\n"
        : "=l"(old.k0), "=l"(old.k1)
        : "l"(0ull), "l"(0ull), "l"(desired.k0), "l"(desired.k1), "l"(slot)
        : "memory");
    return old;
}

__device__ inline void gtable_add(GEntry *gt, uint32_t gmask, const uint8_t *bytes, const Key &key,
                                  uint64_t n, unsigned int *err, uint32_t op = OP_SUM)
{
    uint32_t slot = hash_key(key) & gmask;
    for (uint32_t probes = 0; probes < GT_MAX_PROBES; ++probes) {
        ulonglong2 v = __ldcg((const ulonglong2 *)&gt[slot]);
        Key cur;
        cur.k0 = v.x; cur.k1 = v.y;
        if (cur.k1 == 0) {
            cur = cas128(&gt[slot], key);
            if (cur.k1 == 0) cur = key; // the slot is ours
        }
        if (same_key(cur, key, bytes)) {
            op_apply(&gt[slot].cnt, (unsigned long long)n, op);
            return;
        }
        slot = (slot + 1) & gmask;
    }
    atomicOr(err, ERR_TABLE_FULL);
}

// Length (up to the next whitespace byte) and hash of the token starting at off.
// Returns 0 if the token is longer than max_len.
__host__ __device__ inline uint32_t scan_long_token(const uint8_t *bytes, uint64_t off, uint32_t max_len,
                                           uint32_t *hash_out)
{
    uint64_t h = 0xCBF29CE484222325ull;
    uint32_t len = 0;
    for (;;) {
        uint64_t w = load8(bytes, off + len);
        uint32_t f0 = space_flags((uint32_t)w), f1 = space_flags((uint32_t)(w >> 32));
        uint32_t nb = 8;
        if (f0) nb = (uint32_t)(ffs32(f0) - 8) >> 3;
        else if (f1) nb = 4 + ((uint32_t)(ffs32(f1) - 8) >> 3);
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
