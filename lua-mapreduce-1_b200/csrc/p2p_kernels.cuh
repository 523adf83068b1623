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
//                                           [ records: GEntry 32 B {k0, k1, count, partition} ]
//                                           [ bytes of keys longer than 15 bytes ]
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

// 32-byte load that takes part in the memory model (ld256 is non-coherent: fine for data written
// by earlier kernels, not for records a peer stored while this kernel was already running)
__device__ __forceinline__ ulonglong4 ld256_coherent(const void *p)
{
    ulonglong4 r;
    asm volatile("ld.global.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
                 : "l"(p) : "memory");
    return r;
}

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// entries -> records in the owners' slots.  Each CTA takes one contiguous chunk of entries: tally
// per owner, reserve ranges with one atomic per owner, hand out places from shared-memory cursors.
// The last CTA to finish publishes this rank's epoch in every owner's flag array (release at
// system scope after all CTAs fenced their remote stores).
__global__ void __launch_bounds__(256) p2p_scatter_kernel(P2PParams p, EntryArrays ent, const uint8_t *bytes,
                                                          const unsigned long long *nrec_total,
                                                          const unsigned long long *nbytes_total,
                                                          unsigned long long *rec_cur, unsigned long long *byte_cur,
                                                          unsigned int *done_ctas, unsigned long long max_keys, Ctl *ctl)
{
    __shared__ unsigned int s_rec[P2P_MAX_WORLD];
    __shared__ unsigned long long s_bytes[P2P_MAX_WORLD], s_rec_base[P2P_MAX_WORLD], s_byte_base[P2P_MAX_WORLD];
    __shared__ bool s_last;
    __shared__ unsigned int s_full;
    // a segment that does not fit its slot: nothing is written but poisoned headers, so that sender
    // and owners all report it; every CTA works that out for itself from the tallies
    if (threadIdx.x == 0) s_full = 0;
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x)
        if (P2P_HDR + nrec_total[d] * P2P_REC + nbytes_total[d] > p.slot_bytes) s_full = 1;
    __syncthreads();
    const bool full = s_full;
    if (blockIdx.x == 0) {
        if (full && threadIdx.x == 0) atomicOr(&ctl->err, ERR_SHUFFLE_FULL);
        for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
            ulonglong2 *hdr = (ulonglong2 *)(p.peer_buf[d] + (unsigned long long)p.rank * p.slot_bytes);
            hdr[0] = full ? make_ulonglong2(~0ull, 0) : make_ulonglong2(nrec_total[d], nbytes_total[d]);
            hdr[1] = make_ulonglong2(0, 0);
        }
    }
    const uint64_t n = ctl->n < max_keys ? ctl->n : max_keys;
    const uint64_t chunk = (n + gridDim.x - 1) / gridDim.x;
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    if (!full && lo < hi) {
        for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) { s_rec[d] = 0; s_bytes[d] = 0; }
        __syncthreads();
        for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
            const uint32_t d = ent.owner(i);
            atomicAdd(&s_rec[d], 1u);
            const Key k = ent.key(i);
            if (key_is_long(k)) atomicAdd(&s_bytes[d], (unsigned long long)key_len(k));
        }
        __syncthreads();
        for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
            s_rec_base[d] = s_rec[d] ? atomicAdd(&rec_cur[d], (unsigned long long)s_rec[d]) : 0;
            s_byte_base[d] = s_bytes[d] ? atomicAdd(&byte_cur[d], s_bytes[d]) : 0;
            s_rec[d] = 0; s_bytes[d] = 0;
        }
        __syncthreads();
        for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
            const ulonglong4 e = ld256(&ent.e[i]);
            Key k; k.k0 = e.x; k.k1 = e.y;
            const uint32_t d = (uint32_t)(e.w >> 32);
            const unsigned long long slot_at = (unsigned long long)p.rank * p.slot_bytes; // offset in the owner's buffer
            uint8_t *slot = p.peer_buf[d] + slot_at;
            const unsigned long long r = s_rec_base[d] + atomicAdd(&s_rec[d], 1u);
            if (key_is_long(k)) {
                const uint32_t len = key_len(k);
                const unsigned long long bo = P2P_HDR + nrec_total[d] * P2P_REC + s_byte_base[d] +
                                              atomicAdd(&s_bytes[d], (unsigned long long)len);
                uint8_t *dst = slot + bo;
                const uint64_t off = key_off(k);
                for (uint32_t j = 0; j < len; ++j) dst[j] = bytes[off + j];
                k = make_long_key(slot_at + bo, len, (uint32_t)k.k1); // offset as the owner will see it
            }
            ulonglong2 *rec = (ulonglong2 *)(slot + P2P_HDR + r * P2P_REC);
            rec[0] = make_ulonglong2(k.k0, k.k1);
            rec[1] = make_ulonglong2(e.z, e.w);
        }
    }
    __threadfence_system(); // this thread's remote stores are performed before the CTA is counted as done
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(done_ctas, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence_system();
    for (uint32_t d = threadIdx.x; d < p.world; d += blockDim.x) {
        unsigned long long *flag = (unsigned long long *)p.peer_win[d] + p.rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(p.epoch) : "memory");
    }
    if (threadIdx.x == 0) *done_ctas = 0; // ready for the next shuffle (stream order protects it)
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
    uint32_t hi_or = 0, hi_and = ~0u;
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
            const uint64_t ck = sort_chunk(k, buf, 0);
            SortItem it; it.lo = ck; it.hi = part; it.val = (uint32_t)(ent0 + first + i);
            items[base + i] = it;
            hi_or |= part; hi_and &= part; lo_or |= ck; lo_and &= ck;
            any = true;
        }
    }
    orand_commit(ctl, hi_or, hi_and, lo_or, lo_and, any);
}

} // namespace mr
