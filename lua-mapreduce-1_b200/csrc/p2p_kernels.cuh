// p2p_kernels.cuh -- the shuffle with no staging buffer and no collective call: every rank writes its
// combined (key, count) records straight into the owner's receive window over NVLink (peer memory
// mapped into this process), raises a flag there, and the owner indexes its window in place.
//
// Replaces, for ranks of one NVLink domain, the run files of job.lua:208-221 (write) and
// job.lua:253-260 (read).  Window layout on every rank (one allocation, exported to the peers):
//
//   [ flags: P2P_FLAG_BYTES ][ buffer 0: world slots ][ buffer 1: world slots ]
//
//   slot s of a buffer = what rank s sent:  [ header 32 B {records, key bytes, 0, 0} ]
//                                           [ records: GEntry 32 B {k0, k1, count, partition} ... ->
//                                           <- ... bytes of keys longer than 15 bytes ] (slot end)
//
// Records are table entries already (same 32-byte layout), long keys carry offsets relative to
// the start of the receiving buffer, so the owner sorts and reduces them where they lie.
// Buffers alternate by epoch (a counter every rank advances once per shuffle); flags hold the
// last epoch each source finished writing.  Why two buffers are enough: a sender reaches the
// scatter of epoch e+2 only after its own wait of epoch e+1, hence after the owner signalled
// e+1, which the owner's stream does after it reduced epoch e.
#pragma once
#include "common.cuh"
#include "reduce_kernels.cuh"

namespace mr {

constexpr uint32_t P2P_FLAG_BYTES = 8192;      // world_size <= 1024 flags of 8 bytes
constexpr uint32_t P2P_HDR = 32;
constexpr uint32_t P2P_REC = 32;
constexpr uint32_t P2P_MAX_CHUNKS = 1024;      // index kernel: CTAs per source slot (chosen so that the grid fills the SMs)
constexpr uint32_t P2P_MAX_WORLD = 1024;
constexpr uint32_t ERR_PEER_TIMEOUT = 16u;     // joins the ERR_* bits of common.cuh
constexpr unsigned long long P2P_TIMEOUT_NS = 20ull * 1000 * 1000 * 1000;

struct P2PParams {
    uint8_t *const *peer_buf;   // [world] device array: peer d's buffer of this epoch, as mapped here
    uint8_t *const *peer_win;   // [world] device array: peer d's window (flags at the front)
    uint32_t world, rank;
    unsigned long long slot_bytes;
    unsigned long long epoch;
};

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// The sender's side in ONE kernel: global table -> records in the owners' slots.  A tile of table
// slots is packed in shared memory (so that the partitioner runs on full warps), every record
// gets its owner, the tile reserves its places in each owner's slot with one atomic per owner,
// and the records go straight out over NVLink -- no dense entry array, no second pass.  Records
// grow upwards from the slot header, bytes of keys longer than 15 bytes downwards from the slot
// end, so no total has to be known before the first store.  The last CTA to finish writes the
// headers (poisoned if records and bytes of some slot met: sender and owner both report it) and
// publishes this rank's epoch in every owner's flag array (release at system scope, after every
// CTA fenced its remote stores).
constexpr int P2P_SEND_ITEMS = 4;
__global__ void __launch_bounds__(256) p2p_send_kernel(P2PParams p, GEntry *gt, uint64_t nslots,
                                                       const uint8_t *bytes, uint32_t num_reducers,
                                                       uint32_t partition_kind, OwnerMap owners,
                                                       unsigned long long *rec_cur, unsigned long long *byte_cur,
                                                       unsigned int *done_ctas, Ctl *ctl)
{
    constexpr int ITEMS = P2P_SEND_ITEMS;
    __shared__ ulonglong4 s_ent[256 * ITEMS];
    __shared__ unsigned short s_rl[256 * ITEMS];         // place of the record among the tile's records for its owner
    __shared__ unsigned int s_bl[256 * ITEMS];           // same for its key bytes
    __shared__ unsigned int wsum[8];
    __shared__ bool s_last;
    // per-owner tallies and reserved bases: 24 bytes per rank of dynamic shared memory
    extern __shared__ __align__(16) unsigned char p2p_dyn[];
    unsigned long long *s_rec_base = (unsigned long long *)p2p_dyn, *s_byt_base = s_rec_base + p.world;
    unsigned int *s_rec = (unsigned int *)(s_byt_base + p.world), *s_byt = s_rec + p.world;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t tile_slots = 256 * ITEMS;
    const uint64_t ntiles = (nslots + tile_slots - 1) / tile_slots;
    const unsigned long long slot_at = (unsigned long long)p.rank * p.slot_bytes; // offset in the owner's buffer
    unsigned long long sent = 0;
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        ulonglong4 e[ITEMS];
        uint32_t nocc = 0;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint64_t slot = tile * tile_slots + (uint64_t)i * 256 + threadIdx.x;
            e[i] = make_ulonglong4(0, 0, 0, 0);
            if (slot < nslots) e[i] = ld256_coherent(&gt[slot]);
            if (e[i].y != 0) {
                ((ulonglong4 *)gt)[slot] = make_ulonglong4(0, 0, 0, 0); // the table is left empty for the next job
                ++nocc;
            }
        }
        for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) { s_rec[d] = 0; s_byt[d] = 0; }
        uint32_t incl = nocc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        uint32_t woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { uint32_t c = wsum[w]; if (w < (int)warp) woff += c; total += c; }
        uint32_t pos = woff + incl - nocc;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (e[i].y != 0) s_ent[pos++] = e[i];
        __syncthreads();
        // owner and place inside the tile, per record
        for (uint32_t t = threadIdx.x; t < total; t += 256) {
            ulonglong4 v = s_ent[t];
            Key key; key.k0 = v.x; key.k1 = v.y;
            const uint32_t part = partition_kind == 1 ? __byte_perm((uint32_t)key.k0, 0, 0x0123) % num_reducers
                                                      : partition_of_key(key, bytes, num_reducers);
            const uint32_t d = owners.world > 1 ? owner_of(owners, part, sort_chunk(key, bytes, 0)) : 0;
            s_ent[t].w = part | (unsigned long long)d << 32;
            s_rl[t] = (unsigned short)atomicAdd(&s_rec[d], 1u);
            if (key_is_long(key)) s_bl[t] = atomicAdd(&s_byt[d], key_len(key));
        }
        __syncthreads();
        for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
            s_rec_base[d] = s_rec[d] ? atomicAdd(&rec_cur[d], (unsigned long long)s_rec[d]) : 0;
            s_byt_base[d] = s_byt[d] ? atomicAdd(&byte_cur[d], (unsigned long long)s_byt[d]) : 0;
        }
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < total; t += 256) {
            const ulonglong4 v = s_ent[t];
            Key k; k.k0 = v.x; k.k1 = v.y;
            const uint32_t d = (uint32_t)(v.w >> 32);
            uint8_t *slot = p.peer_buf[d] + slot_at;
            const unsigned long long rec_end = P2P_HDR + (s_rec_base[d] + s_rl[t] + 1) * P2P_REC;
            bool fits = rec_end <= p.slot_bytes;
            if (key_is_long(k)) {
                const uint32_t len = key_len(k);
                const unsigned long long below = s_byt_base[d] + s_bl[t] + len; // bytes from the slot end down to this key's start
                fits = fits && below + P2P_HDR <= p.slot_bytes;
                if (fits) {
                    const unsigned long long bo = p.slot_bytes - below;
                    uint8_t *dst = slot + bo;
                    const uint64_t off = key_off(k);
                    for (uint32_t j = 0; j < len; ++j) dst[j] = bytes[off + j];
                    k = make_long_key(slot_at + bo, len, (uint32_t)k.k1); // offset as the owner will see it
                }
            }
            if (fits) { // a record that does not fit is simply not stored: the header check below reports the slot
                ulonglong2 *rec = (ulonglong2 *)(slot + rec_end - P2P_REC);
                rec[0] = make_ulonglong2(k.k0, k.k1);
                rec[1] = make_ulonglong2(v.z, v.w & 0xFFFFFFFFull);
            }
        }
        sent += total;
        __syncthreads();
    }
    if (threadIdx.x == 0 && sent) atomicAdd(&ctl->sent, sent); // records this rank sent (mr_job_info.map_keys)
    // the CTA's remote stores are ordered before it is counted as done: the barrier makes them
    // visible to thread 0, whose system-scope fence is cumulative over them (one fence per CTA --
    // a fence in every thread cost a third of this kernel's time)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(done_ctas, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
        const unsigned long long nr = *(volatile unsigned long long *)&rec_cur[d], nb = *(volatile unsigned long long *)&byte_cur[d];
        const bool full = P2P_HDR + nr * P2P_REC + nb > p.slot_bytes;
        if (full) atomicOr(&ctl->err, ERR_SHUFFLE_FULL);
        ulonglong2 *hdr = (ulonglong2 *)(p.peer_buf[d] + slot_at);
        hdr[0] = full ? make_ulonglong2(~0ull, 0) : make_ulonglong2(nr, nb);
        hdr[1] = make_ulonglong2(0, 0);
    }
    __threadfence_system();
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
        unsigned long long *flag = (unsigned long long *)p.peer_win[d] + p.rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(p.epoch) : "memory");
    }
    if (threadIdx.x == 0) *done_ctas = 0; // ready for the next shuffle (stream order protects it)
}

// 8 bytes at an arbitrary offset with loads that take part in the memory model: the bytes were
// stored by a peer while this kernel may already have been running, so the read-only (.nc) path of
// load8 is not allowed for them
__device__ __forceinline__ uint64_t load8_coherent(const uint8_t *base, uint64_t off)
{
    const volatile unsigned long long *p = (const volatile unsigned long long *)(base + (off & ~7ull));
    const uint32_t sh = (uint32_t)(off & 7) * 8;
    const uint64_t w0 = p[0];
    if (sh == 0) return w0;
    const uint64_t w1 = p[1];
    return (w0 >> sh) | (w1 << (64 - sh));
}

// The owner's side: wait for source src (one CTA group per source, so that early arrivals are
// indexed while late ones are still in flight), then turn its records into sort items in place.
// Item order is whatever the atomics give; the sort does not care.
__global__ void __launch_bounds__(256) p2p_index_kernel(const uint8_t *buf, const unsigned long long *flags,
                                                        uint32_t world, uint32_t chunks, unsigned long long slot_bytes,
                                                        unsigned long long epoch, SortItem *items,
                                                        unsigned long long max_keys, Ctl *ctl)
{
    __shared__ unsigned long long s_n, s_base;
    __shared__ bool s_ok;
    __shared__ unsigned int s_occ[56];
    occ_clear(s_occ);
    const uint32_t src = blockIdx.x / chunks, chunk = blockIdx.x % chunks;
    const unsigned long long slot_at = (unsigned long long)src * slot_bytes;
    if (threadIdx.x == 0) {
        const unsigned long long t0 = globaltimer_ns();
        unsigned long long seen;
        bool ok = true;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flags + src) : "memory");
            if (seen >= epoch) break;
            if (globaltimer_ns() - t0 > P2P_TIMEOUT_NS) { ok = false; break; }
            __nanosleep(200);
        }
        unsigned long long n = 0;
        if (!ok) {
            atomicOr(&ctl->err, ERR_PEER_TIMEOUT);
        } else {
            const unsigned long long *hdr = (const unsigned long long *)(buf + slot_at);
            n = hdr[0];
            if (P2P_HDR + n * P2P_REC + hdr[1] > slot_bytes || n > slot_bytes) { n = 0; atomicOr(&ctl->err, ERR_SHUFFLE_FULL); }
        }
        const unsigned long long per = (n + chunks - 1) / chunks;
        const unsigned long long lo = per * chunk, hi = lo + per < n ? lo + per : n;
        const unsigned long long mine = hi > lo ? hi - lo : 0;
        unsigned long long base = mine ? atomicAdd(&ctl->n, mine) : 0;
        s_n = mine; s_base = base; s_ok = ok;
    }
    __syncthreads();
    uint32_t hi_or = 0, hi_and = ~0u, pmin = ~0u, pmax = 0;
    uint64_t lo_or = 0, lo_and = ~0ull;
    bool any = false;
    if (s_ok && s_n) {
        const unsigned long long n_here = s_n, base = s_base;
        const unsigned long long *hdr = (const unsigned long long *)(buf + slot_at);
        const unsigned long long n = hdr[0];
        const unsigned long long per = (n + chunks - 1) / chunks;
        const unsigned long long first = per * chunk;
        const GEntry *rec = (const GEntry *)(buf + slot_at + P2P_HDR);
        const unsigned long long ent0 = (slot_at + P2P_HDR) / P2P_REC; // entry index of record 0 (slot_bytes % 32 == 0)
        for (unsigned long long i = threadIdx.x; i < n_here; i += blockDim.x) {
            if (base + i >= max_keys) { atomicOr(&ctl->err, ERR_KEYS_FULL); break; }
            const ulonglong4 e = ld256_coherent(&rec[first + i]);
            Key k; k.k0 = e.x; k.k1 = e.y;
            const uint32_t part = (uint32_t)e.w;
            uint64_t ck;
            if (key_is_long(k)) {
                // the record names its key bytes by offset: they must lie inside the sender's slot
                // (a poisoned record is reported, never followed), and are read coherently
                const uint64_t off = key_off(k), len = key_len(k);
                if (off < slot_at + P2P_HDR || off + len > slot_at + slot_bytes) {
                    atomicOr(&ctl->err, ERR_SHUFFLE_FULL);
                    ck = 8;
                } else {
                    ck = (bswap64(load8_coherent(buf, off) & low_bytes_mask(7)) & ~0xFFull) | 8; // = sort_chunk(k, buf, 0)
                }
            } else {
                ck = sort_chunk(k, buf, 0);
            }
            SortItem it; it.lo = ck; it.hi = part; it.val = (uint32_t)(ent0 + first + i);
            items[base + i] = it;
            hi_or |= part; hi_and &= part; lo_or |= ck; lo_and &= ck;
            pmin = min(pmin, part); pmax = max(pmax, part);
            occ_add(s_occ, ck);
            any = true;
        }
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any, s_occ, pmin, pmax);
}

} // namespace mr
