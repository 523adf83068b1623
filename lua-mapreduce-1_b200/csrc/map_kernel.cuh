// map_kernel.cuh -- K1+K2: tokenise raw split bytes and hash-combine (token -> count).
//
// Replaces, for MR_MAP_TOKENIZE_WS + MR_REDUCE_SUM_U64, the reference's per-token Lua
// path: mapfn.lua:3-9 (io.lines + gmatch("[^%s]+")), emit (job.lua:83-97) and the
// combiner (reducefn.lua:1-5 through job.lua:92-96,198-202).
//
// Layout.  The arena holds the staged splits back to back, each starting on a
// 16-byte boundary, every gap and the tail filled with 0x20, so split boundaries are
// token boundaries and the kernel never needs to know where splits start.  The arena
// is cut into segments of SEG = 1024 bytes; a warp owns a segment at a time:
//   1. each lane loads its 32 consecutive bytes (2 x 16-byte streaming loads, one 32-byte
//      sector per lane), writes them to the warp's staging buffer and derives a 32-bit
//      non-whitespace mask with byte-parallel arithmetic; the 32 masks form a 1024-bit
//      bitmap in shared memory (plus one word for the 32 look-ahead bytes);
//   2. token starts (mask & ~(mask << 1 | carry-from-previous-lane)) are counted,
//      warp-scanned and their positions compacted into a per-warp list, so that
//   3. every lane takes ONE token: its length from the bitmap (funnel shift +
//      find-first-zero), 16 bytes from the staging buffer (5 aligned words + funnel
//      shifts) packed into the 128-bit inline key.  Keys of <= 7 bytes go to the CTA's
//      shared-memory table: 4-way buckets read with two 16-byte loads, no probe loop,
//      ATOMS.ADD on a hit, ATOMS.CAS.64 only on first sight.
// Keys the CTA table does not hold (bucket full, or 8..15 bytes) are queued per warp and
// inserted into the global table 32 at a time, so global-memory latency is paid by full
// warps and never inside the per-token loop; tokens of 16+ bytes take a separate rare
// path that keeps a reference to their bytes.  At the end every CTA flushes its table
// into the global one.
#pragma once
#include "common.cuh"

// Measured variants (scripts/map_variants.py, profiles/r02/map_variants_*.txt).  On by default: full
// batches of medium tokens go from registers straight into flight (-1.1 %); -DMRGPU_V_QUEUEMED restores
// the trip through the per-warp queue.
#if !defined(MRGPU_V_QUEUEMED) && !defined(MRGPU_V_DIRECTMED)
#define MRGPU_V_DIRECTMED 1
#endif

namespace mr {

constexpr int SEG = 1024;              // bytes per warp segment
constexpr int SEG_AHEAD = 32;          // look-ahead bytes staged past the segment
constexpr int WAYS = 4;

struct MapParams {
    const uint8_t *arena;
    uint64_t first_seg, nseg; // segments [first_seg, first_seg + nseg)
    GEntry *gtab;
    uint32_t gmask;
    uint32_t max_token;
    unsigned long long *tokens; // running token count
    unsigned int *err;
    long long *dbg;             // MRGPU_MAP_TIMING builds: per-phase cycles of one warp
};

#ifdef MRGPU_MAP_TIMING
#define MT_DECL long long mt_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, mt_t = clock64(), mt_n
#define MT(k) do { mt_n = clock64(); mt_acc[k] += mt_n - mt_t; mt_t = mt_n; } while (0)
#define MT_END do { if (blockIdx.x == 7 && warp == 5 && lane == 0) for (int q = 0; q < 8; ++q) p.dbg[q] = mt_acc[q]; } while (0)
#else
#define MT_DECL do { } while (0)
#define MT(k) do { } while (0)
#define MT_END do { } while (0)
#endif

// per-warp shared memory
struct __align__(16) WarpSmem {
    uint32_t stage[(SEG + SEG_AHEAD) / 4]; // 1056 B: the segment + 32 look-ahead bytes
    uint16_t list[SEG / 2];                // token entries (pos | len << 10): short tokens from the
                                           // front, medium tokens from the back (max one per 2 bytes)
    Key gq[63];                            // keys waiting for the global table: <= 31 after a drain, + 32 per iteration
    unsigned long long longq[8];           // arena offsets of 16+ byte tokens
};
// The size of this block decides the kernel's speed in an unexpected way: 32 of them + the 96 KB table
// must fit the SM's 196 KB shared-memory configuration, because the next one (228 KB) leaves 28 KB of
// L1 -- too little for the loads a CTA of 32 warps keeps in flight (input sectors + one batch of table
// probes per warp).  Measured on config 2: 525 us in the 228 KB configuration, 446 us in the 196 KB one
// (profiles/r02/map_variants_3..5.txt, map_carveout*.txt; the same kernel forced back with
// cudaFuncAttributePreferredSharedMemoryCarveout = 100 is slow again).
static_assert(sizeof(WarpSmem) == 3152, "32 warps + the 96 KB table must fit the 196 KB shared-memory configuration");
constexpr uint32_t GQ_KEEP = 32, LONGQ_DRAIN = 6;

template <int LOG2_SLOTS>
struct MapSmemLayout {
    static constexpr int SLOTS = 1 << LOG2_SLOTS;
    static constexpr size_t keys_bytes = (size_t)SLOTS * 8;
    static constexpr size_t cnt_bytes = (size_t)SLOTS * 4;
    static constexpr size_t warp_off = keys_bytes + cnt_bytes;
    static constexpr size_t total(int warps) { return warp_off + (size_t)warps * sizeof(WarpSmem); }
};

__device__ __forceinline__ uint4 ld_stream(const uint4 *p)
{
    // streaming read: the input is touched once, keep L2 for the hash table
    uint4 r;
    asm volatile("ld.global.cs.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ ulonglong2 lds128(uint32_t saddr)
{
    // one 16-byte shared-memory load of two table keys (other warps update the table concurrently)
    ulonglong2 r;
#ifdef MRGPU_V_PLAINLDS
    asm("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "r"(saddr));
#else
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "r"(saddr));
#endif
    return r;
}

// slot of (bucket, way) in the split layout: [ways 0-1 of all buckets][ways 2-3 of all buckets]
__device__ __forceinline__ uint32_t slot_of(uint32_t bucket, int way, uint32_t slots)
{
    return (way >> 1) * (slots / 2) + bucket * 2 + (way & 1);
}

__device__ __forceinline__ void drain_keys(const MapParams &p, WarpSmem *ws, uint32_t first, uint32_t n,
                                           uint32_t lane)
{
    if (lane < n) {
        const ulonglong2 v = *(const ulonglong2 *)&ws->gq[first + lane]; // one 16-byte load
        Key k;
        k.k0 = v.x; k.k1 = v.y & ~(0xFull << 60); // a re-queued key still carries its probe distance
        gtable_add(p.gtab, p.gmask, p.arena, k, 1, p.err);
    }
}

// Asynchronous insertion into the global table.  A queued key carries its probe distance in
// bits 60..63 of k1 (free for inline keys, whose length uses bits 56..59).  probe_issue starts the
// 16-byte load of slot (home + distance); probe_try, called a few hundred instructions later, finishes
// the insertion if that slot holds the key or is empty, and otherwise reports "try the next slot",
// upon which the caller simply re-queues the key with distance + 1: no drain ever waits for memory.
// A stale loaded value is harmless: an empty slot is claimed with compare-and-swap, which returns
// the current content, and a slot's key never changes once set.
constexpr unsigned long long QDIST_MASK = 0xFull << 60;

__device__ __forceinline__ void probe_issue(const MapParams &p, const Key &qkey, uint32_t &slot, ulonglong2 &v)
{
    Key key = qkey;
    key.k1 &= ~QDIST_MASK;
    slot = (hash_key(key) + (uint32_t)(qkey.k1 >> 60)) & p.gmask;
    v = __ldcg((const ulonglong2 *)&p.gtab[slot]);
}

#ifdef MRGPU_V_ASYNCCLAIM
// an empty slot is claimed WITHOUT waiting for the compare-and-swap's answer (a round trip to L2 with the
// whole warp stalled): the key is probed again at the same slot on the next round, where it finds
// itself or the key that won the race
__device__ __forceinline__ void cas128_noret(GEntry *slot, const Key &desired)
{
    asm volatile(
        "{\n\t"
        ".reg .b128 cmp, des, res;\n\t"
        "mov.b128 cmp, {%0, %1};\n\t"
        "mov.b128 des, {%2, %3};\n\t"
        "atom.global.cas.b128 res, [%4], cmp, des;\n\t"
        "}\n"
        :: "l"(0ull), "l"(0ull), "l"(desired.k0), "l"(desired.k1), "l"(slot)
        : "memory");
}
#endif

// returns true when the key has been counted; *same_slot (optional): probe this slot again
__device__ __forceinline__ bool probe_try(const MapParams &p, const Key &qkey, uint32_t slot, ulonglong2 v, bool *same_slot = nullptr)
{
    Key key = qkey, cur;
    key.k1 &= ~QDIST_MASK;
    cur.k0 = v.x; cur.k1 = v.y;
    if (cur.k1 == 0) {
#ifdef MRGPU_V_ASYNCCLAIM
        if (same_slot) { cas128_noret(&p.gtab[slot], key); *same_slot = true; return false; }
#endif
        cur = cas128(&p.gtab[slot], key);
        if (cur.k1 == 0) cur = key; // the slot is ours
    }
    if (cur.k0 == key.k0 && cur.k1 == key.k1) {
        atomicAdd(&p.gtab[slot].cnt, 1ull);
        return true;
    }
    return false;
}

__device__ __forceinline__ void drain_long(const MapParams &p, WarpSmem *ws, uint32_t n, uint32_t lane)
{
    if (lane < n) {
        uint64_t off = ws->longq[lane];
        uint32_t h;
        uint32_t len = scan_long_token(p.arena, off, p.max_token, &h);
        if (len == 0) atomicOr(p.err, ERR_TOKEN_TOO_LONG);
        else gtable_add(p.gtab, p.gmask, p.arena, make_long_key(off, len, h), 1, p.err);
    }
}

#ifndef MRGPU_V_CTAS
#define MRGPU_V_CTAS 1
#endif
template <int WARPS, int LOG2_SLOTS>
__global__ void __launch_bounds__(WARPS * 32, WARPS < 32 ? MRGPU_V_CTAS : 1) map_combine_kernel(MapParams p)
{
    using L = MapSmemLayout<LOG2_SLOTS>;
    constexpr uint32_t SLOTS = L::SLOTS;
    constexpr uint32_t LIST_LAST = SEG / 2 - 1;
    extern __shared__ __align__(16) uint8_t smem[];
    unsigned long long *tkeys = (unsigned long long *)smem;
    unsigned int *tcnt = (unsigned int *)(smem + L::keys_bytes);
    __shared__ unsigned int tused;                       // keys held by the CTA table
#ifdef MRGPU_V_ONEBYTE
    __shared__ unsigned int c1[256];                     // tokens of one byte: counted by their byte value, no key, no probe
    for (uint32_t i = threadIdx.x; i < 256; i += WARPS * 32) c1[i] = 0;
#endif
    constexpr uint32_t FILL_LIMIT = SLOTS - SLOTS / 8;   // stop trying to insert at 87.5 % load
    const uint32_t tkeys_s = (uint32_t)__cvta_generic_to_shared(tkeys);

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    WarpSmem *ws = (WarpSmem *)(smem + L::warp_off) + warp;

#ifdef MRGPU_MAP_TIMING
    const long long mt_k0 = clock64();
#endif
    for (uint32_t i = tid; i < SLOTS; i += WARPS * 32) { tkeys[i] = 0; tcnt[i] = 0; }
    if (tid == 0) tused = 0;
    __syncthreads();
#ifdef MRGPU_MAP_TIMING
    if (blockIdx.x == 7 && warp == 5 && lane == 0) p.dbg[8] = clock64() - mt_k0;
#endif

    const uint32_t lt = lanemask_lt();
    uint32_t nq = 0, nlong = 0; // warp-uniform fills of gq / longq
    // one batch of 32 keys whose slot loads are in flight (lane i holds key i)
    bool pending = false;
    Key pkey;
    pkey.k0 = 0; pkey.k1 = 0;
    uint32_t pslot = 0;
    ulonglong2 pv = make_ulonglong2(0, 0);
    // take the newest 32 queued keys, finish the batch in flight (re-queueing what needs another
    // probe), then put the new batch in flight
    auto drain32 = [&]() {
        nq -= 32;
        const ulonglong2 kv = *(const ulonglong2 *)&ws->gq[nq + lane];
        __syncwarp();
        if (pending) {
            bool same = false;
            bool again = !probe_try(p, pkey, pslot, pv, &same);
            if (again && !same && (pkey.k1 >> 60) == 15) { // very long probe chain: finish it synchronously
                Key k = pkey;
                k.k1 &= ~QDIST_MASK;
                gtable_add(p.gtab, p.gmask, p.arena, k, 1, p.err);
                again = false;
            }
            const uint32_t am = __ballot_sync(0xFFFFFFFFu, again);
            if (again) *(ulonglong2 *)&ws->gq[nq + __popc(am & lt)] = make_ulonglong2(pkey.k0, pkey.k1 + (same ? 0ull : 1ull << 60));
            nq += __popc(am);
        }
        pkey.k0 = kv.x; pkey.k1 = kv.y;
        probe_issue(p, pkey, pslot, pv);
        pending = true;
        __syncwarp();
    };
#ifdef MRGPU_V_DIRECTMED
    // a full batch of 32 keys goes from registers straight into flight (no trip through the queue)
    auto drain_direct = [&](const Key &nk) {
        if (pending) {
            bool again = !probe_try(p, pkey, pslot, pv);
            if (again && (pkey.k1 >> 60) == 15) {
                Key k = pkey;
                k.k1 &= ~QDIST_MASK;
                gtable_add(p.gtab, p.gmask, p.arena, k, 1, p.err);
                again = false;
            }
            const uint32_t am = __ballot_sync(0xFFFFFFFFu, again);
            if (again) *(ulonglong2 *)&ws->gq[nq + __popc(am & lt)] = make_ulonglong2(pkey.k0, pkey.k1 + (1ull << 60));
            nq += __popc(am);
        }
        pkey = nk;
        probe_issue(p, pkey, pslot, pv);
        pending = true;
        __syncwarp();
    };
#endif
    unsigned long long my_tokens = 0;
    const uint64_t gwarp = (uint64_t)blockIdx.x * WARPS + warp, nwarps = (uint64_t)gridDim.x * WARPS;
    MT_DECL;
    for (uint64_t s = gwarp; s < p.nseg; s += nwarps) {
        MT(7);
        const uint64_t base = (p.first_seg + s) * SEG;
        const uint4 *src = (const uint4 *)(p.arena + base);
        // ---- 1. load + stage + non-space mask of the lane's 32 bytes
        const uint4 a = ld_stream(src + 2 * lane), b = ld_stream(src + 2 * lane + 1);
        uint4 ahead = make_uint4(0, 0, 0, 0);
        if (lane < SEG_AHEAD / 16) ahead = __ldg(src + SEG / 16 + lane);
        uint32_t prev_ns = 0; // is the byte before the segment a token byte?
        if (base != 0) {
            const uint32_t c = __ldg(p.arena + base - 1);
            prev_ns = !(c == 0x20u || (c - 9u) < 5u);
        }
        uint4 *st4 = (uint4 *)ws->stage;
        st4[2 * lane] = a;
        st4[2 * lane + 1] = b;
        if (lane < SEG_AHEAD / 16) st4[SEG / 16 + lane] = ahead;
        const uint32_t m = nonspace_nibble(a.x) | nonspace_nibble(a.y) << 4 | nonspace_nibble(a.z) << 8 |
                           nonspace_nibble(a.w) << 12 | nonspace_nibble(b.x) << 16 |
                           nonspace_nibble(b.y) << 20 | nonspace_nibble(b.z) << 24 |
                           nonspace_nibble(b.w) << 28;
        __syncwarp();
        MT(0); // loads + staging + mask
        // mask of the 32 look-ahead bytes: lanes 0..7 take one word each
        uint32_t am = 0;
        if (lane < 8) am = nonspace_nibble(ws->stage[SEG / 4 + lane]) << (4 * lane);
        am |= __shfl_xor_sync(0xFFFFFFFFu, am, 1);
        am |= __shfl_xor_sync(0xFFFFFFFFu, am, 2);
        am |= __shfl_xor_sync(0xFFFFFFFFu, am, 4);
        am = __shfl_sync(0xFFFFFFFFu, am, 0);
        // ---- 2. token starts, lengths and classes from the masks
        uint32_t carry = __shfl_up_sync(0xFFFFFFFFu, m >> 31, 1);
        if (lane == 0) carry = prev_ns;
        uint32_t nxt = __shfl_down_sync(0xFFFFFFFFu, m, 1);
        if (lane == 31) nxt = am;
        const TokenMasks tm = token_masks(m, nxt, carry);
        const unsigned long long w = tm.w;
        const uint32_t starts = tm.starts, s_long = tm.s_long, s_med = tm.s_med, s_short = tm.s_short;
#ifdef MRGPU_V_ONEBYTE
        const uint32_t s_one = s_short & ~(uint32_t)(w >> 1);      // a token byte followed by a non-token byte
        const uint32_t s_sh2 = s_short & ~s_one;
        const uint32_t packed = __popc(s_sh2) | (__popc(s_med) << 10) | (__popc(s_one) << 20);
#else
        const uint32_t packed = __popc(s_short) | (__popc(s_med) << 16);
#endif
        uint32_t incl = packed;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        const uint32_t totals = __shfl_sync(0xFFFFFFFFu, incl, 31);
#ifdef MRGPU_V_ONEBYTE
        const uint32_t n_short = totals & 0x3FFu, n_med = (totals >> 10) & 0x3FFu, n_one = totals >> 20;
        uint32_t o_s = (incl - packed) & 0x3FFu;
        uint32_t o_m = LIST_LAST - (((incl - packed) >> 10) & 0x3FFu);
        uint32_t o_1 = n_short + ((incl - packed) >> 20);          // the one-byte tokens follow the short ones in the list
        uint32_t st = starts & ~s_long;
        while (st) {
            const uint32_t bit = __ffs((int)st) - 1;
            st &= st - 1;
            const uint32_t len = token_len(w, bit); // 1..15
            const uint16_t e = (uint16_t)((lane * 32 + bit) | (len << 10));
            if ((s_med >> bit) & 1u) ws->list[o_m--] = e;
            else if (len == 1) ws->list[o_1++] = e;
            else ws->list[o_s++] = e;
        }
#else
        const uint32_t n_short = totals & 0xFFFFu, n_med = totals >> 16;
        uint32_t o_s = (incl - packed) & 0xFFFFu;
        uint32_t o_m = LIST_LAST - ((incl - packed) >> 16);
        uint32_t st = starts & ~s_long;
        while (st) {
            const uint32_t bit = __ffs((int)st) - 1;
            st &= st - 1;
            const uint32_t len = token_len(w, bit); // 1..15
            const uint16_t e = (uint16_t)((lane * 32 + bit) | (len << 10));
            if ((s_med >> bit) & 1u) ws->list[o_m--] = e;
            else ws->list[o_s++] = e;
        }
#endif
        my_tokens += __popc(starts);
        // tokens of 16+ bytes (rare): queue their arena offsets
        uint32_t lb = __ballot_sync(0xFFFFFFFFu, s_long != 0);
        while (lb) {
            const int srcl = __ffs((int)lb) - 1;
            lb &= lb - 1;
            const uint32_t sl = __shfl_sync(0xFFFFFFFFu, s_long, srcl);
            const uint32_t c = __popc(sl);
            if (lane < c) ws->longq[nlong + lane] = base + srcl * 32 + __fns(sl, 0, lane + 1);
            nlong += c;
            __syncwarp();
            if (nlong > LONGQ_DRAIN) { drain_long(p, ws, nlong, lane); nlong = 0; __syncwarp(); }
        }
        const bool can_insert = *(volatile unsigned int *)&tused < FILL_LIMIT; // warp-uniform
        uint32_t inserted = 0;
        __syncwarp();
        MT(1); // starts, classes, scan, lists
        // ---- 3a. short tokens (<= 7 bytes): one per lane, CTA table
        for (uint32_t t0 = 0; t0 < n_short; t0 += 32) {
            const uint32_t j = t0 + lane;
            bool to_global = false;
            Key key;
            key.k0 = 0; key.k1 = 0;
            if (j < n_short) {
                const uint32_t e = ws->list[j];
                const uint32_t pos = e & 1023u, len = e >> 10;
                const uint32_t wi = pos >> 2, bs = (pos & 3) * 8;
                const uint32_t x0 = ws->stage[wi], x1 = ws->stage[wi + 1], x2 = ws->stage[wi + 2];
                const ShortKey skk = short_key(x0, x1, x2, bs, len);
                const uint32_t w0 = skk.w0, skh = skk.w1;
                const unsigned long long sk = (uint64_t)w0 | ((uint64_t)skh << 32); // bytes + length
                const uint32_t bucket = (w0 * 0x9E3779B1u + skh * 0x85EBCA77u) >> (32 - (LOG2_SLOTS - 2));
                // ways 0-1 live in the first half of the key array, ways 2-3 in the second, so the
                // 16-byte first probe of a bucket falls on all eight bank groups, not four
                const uint32_t baddr = tkeys_s + bucket * 16;
                const ulonglong2 q0 = lds128(baddr);
                int way = q0.x == sk ? 0 : q0.y == sk ? 1 : -1;
                if (way < 0) {
                    const ulonglong2 q1 = lds128(baddr + SLOTS * 4);
                    way = q1.x == sk ? 2 : q1.y == sk ? 3 : -1;
                    if (way < 0 && can_insert) {
                        const int e2 = q0.x == 0 ? 0 : q0.y == 0 ? 1 : q1.x == 0 ? 2 : q1.y == 0 ? 3 : -1;
                        if (e2 >= 0) {
                            const unsigned long long old = atomicCAS(&tkeys[slot_of(bucket, e2, SLOTS)], 0ull, sk);
                            if (old == 0) ++inserted;
                            if (old == 0 || old == sk) way = e2;
                        }
                    }
                }
                if (way >= 0) {
                    atomicAdd(&tcnt[slot_of(bucket, way, SLOTS)], 1u);
                } else {
                    to_global = true;
                    key.k0 = sk & LOW56;
                    key.k1 = sk & ~LOW56;
                }
            }
            const uint32_t gm = __ballot_sync(0xFFFFFFFFu, to_global);
            if (to_global) *(ulonglong2 *)&ws->gq[nq + __popc(gm & lt)] = make_ulonglong2(key.k0, key.k1);
            nq += __popc(gm);
            __syncwarp();
            MT(2); // short tokens: key + CTA table
            if (nq >= 32) { drain32(); while (nq >= GQ_KEEP) drain32(); MT(4); }
        }
#ifdef MRGPU_V_ONEBYTE
        // ---- 3a'. tokens of one byte: the byte value is the counter index
        for (uint32_t t0 = 0; t0 < n_one; t0 += 32) {
            const uint32_t j = t0 + lane;
            if (j < n_one) {
                const uint32_t pos = ws->list[n_short + j] & 1023u;
                atomicAdd(&c1[((const uint8_t *)ws->stage)[pos]], 1u);
            }
        }
#endif
        // ---- 3b. medium tokens (8..15 bytes): 128-bit inline key, straight to the global queue
        for (uint32_t t0 = 0; t0 < n_med; t0 += 32) {
            const uint32_t j = t0 + lane;
#ifdef MRGPU_V_DIRECTMED
            if (n_med - t0 >= 32 && nq < GQ_KEEP) {
                const uint32_t e = ws->list[LIST_LAST - j];
                const uint32_t pos = e & 1023u, len = e >> 10;
                const uint32_t wi = pos >> 2, bs = (pos & 3) * 8;
                const Key key = medium_key(ws->stage[wi], ws->stage[wi + 1], ws->stage[wi + 2], ws->stage[wi + 3], ws->stage[wi + 4], bs, len);
                drain_direct(key);
                // the batch that was in flight may have re-queued up to 32 keys: back below GQ_KEEP before
                // anything else is appended (the queue holds 63: <= 31 kept + 32 new)
                while (nq >= GQ_KEEP) drain32();
                continue;
            }
#endif
            if (j < n_med) {
                const uint32_t e = ws->list[LIST_LAST - j];
                const uint32_t pos = e & 1023u, len = e >> 10;
                const uint32_t wi = pos >> 2, bs = (pos & 3) * 8;
                const uint32_t x0 = ws->stage[wi], x1 = ws->stage[wi + 1], x2 = ws->stage[wi + 2],
                               x3 = ws->stage[wi + 3], x4 = ws->stage[wi + 4];
                const Key key = medium_key(x0, x1, x2, x3, x4, bs, len);
                *(ulonglong2 *)&ws->gq[nq + (j - t0)] = make_ulonglong2(key.k0, key.k1);
            }
            nq += min(32u, n_med - t0);
            __syncwarp();
            MT(3); // medium tokens: key build + queue
            if (nq >= 32) { drain32(); while (nq >= GQ_KEEP) drain32(); MT(4); }
        }
        if (inserted) atomicAdd(&tused, inserted);
        __syncwarp();
    }
    if (pending && !probe_try(p, pkey, pslot, pv)) { // the batch in flight, then the rest, synchronously
        Key k = pkey;
        k.k1 &= ~QDIST_MASK;
        gtable_add(p.gtab, p.gmask, p.arena, k, 1, p.err);
    }
    for (uint32_t q0 = 0; q0 < nq; q0 += 32) drain_keys(p, ws, q0, min(32u, nq - q0), lane);
    drain_long(p, ws, nlong, lane);
    MT(5);
    MT_END;
    for (int d = 16; d; d >>= 1) my_tokens += __shfl_xor_sync(0xFFFFFFFFu, my_tokens, d);
    if (lane == 0 && my_tokens) atomicAdd(p.tokens, my_tokens);
#ifdef MRGPU_MAP_TIMING
    const long long mt_k1 = clock64();
#endif
    __syncthreads();
#ifdef MRGPU_MAP_TIMING
    const long long mt_k2 = clock64();
#endif
#ifdef MRGPU_V_ONEBYTE
    for (uint32_t i = tid; i < 256; i += WARPS * 32)
        if (c1[i]) {
            Key k;
            k.k0 = i;
            k.k1 = 1ull << 56;
            gtable_add(p.gtab, p.gmask, p.arena, k, c1[i], p.err);
        }
#endif
    // flush the CTA table
    for (uint32_t i = tid; i < SLOTS; i += WARPS * 32) {
        const unsigned long long sk = tkeys[i];
        if (sk) {
            Key k;
            k.k0 = sk & LOW56;
            k.k1 = sk & ~LOW56;
            gtable_add(p.gtab, p.gmask, p.arena, k, tcnt[i], p.err);
        }
    }
#ifdef MRGPU_MAP_TIMING
    if (blockIdx.x == 7 && warp == 5 && lane == 0) {
        p.dbg[9] = mt_k2 - mt_k1;            // waiting for the CTA's other warps
        p.dbg[10] = clock64() - mt_k2;       // this warp's share of the flush
        p.dbg[11] = tused;
    }
#endif
}

} // namespace mr
