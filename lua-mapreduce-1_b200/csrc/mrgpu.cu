// mrgpu.cu -- context, staging, launch sequencing and the C ABI of include/mrgpu.h.
// Device code lives in map_kernel.cuh (tokenise + combine) and reduce_kernels.cuh
// (partition, radix sort, tie resolution, segmented reduce, rows).  sm_100a only.
#include "mrgpu.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "map_kernel.cuh"
#include "bigram_kernel.cuh"
#include "records_kernel.cuh"
#include "reduce_kernels.cuh"
#include "msd_kernels.cuh"
#include "p2p_kernels.cuh"

using namespace mr;

static_assert(SORT_THREADS >= RADIX_BINS, "one thread per radix bin");

namespace {

std::string g_create_error;

constexpr uint64_t ARENA_SLACK = SEG + SEG_AHEAD + 256;
constexpr uint64_t TAIL_GUARD = SEG_AHEAD + 64; // spaces after the last mapped segment (look-ahead window)
constexpr uint64_t REC_BYTES = 24; // shuffle record {k0, k1, count}

inline uint64_t round_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

// start of a job: the whole control block to zero, the AND masks to all ones (one launch)
__global__ void ctl_job_reset_kernel(Ctl *ctl)
{
    uint32_t *w = (uint32_t *)ctl;
    for (uint32_t i = threadIdx.x; i < sizeof(Ctl) / 4; i += blockDim.x) w[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0) { ctl->and_lo = ~0ull; ctl->and_hi = ~0u; ctl->part_min = ~0u; }
}

__global__ void ctl_init_kernel(Ctl *ctl)
{
    // everything else was zeroed by the memset before this launch
    ctl->and_lo = ~0ull;
    ctl->and_hi = ~0u;
    ctl->part_min = ~0u;
}

__global__ void ctl_set_n_kernel(Ctl *ctl, unsigned long long n) { ctl->n = n; }

// gaps between staged splits (and the tail) are filled with 0x20 by ONE kernel at map time, so
// that the host->device copies of the splits run back to back on the copy engine
struct Gap { unsigned long long off, len; };
constexpr uint32_t MAX_GAPS = 1u << 15;
__global__ void __launch_bounds__(256) fill_gaps_kernel(uint8_t *arena, const Gap *gaps, uint32_t ngaps, uint8_t fill)
{
    for (uint32_t g = blockIdx.x; g < ngaps; g += gridDim.x) {
        const Gap gp = gaps[g];
        for (unsigned long long i = threadIdx.x; i < gp.len; i += blockDim.x) arena[gp.off + i] = fill;
    }
}

// what the owner's side of a shuffle starts from: counters of the reduce side cleared, token
// count, error bits and key pool of the map side kept
__global__ void ctl_owner_reset_kernel(Ctl *ctl)
{
    ctl->n = 0;
    ctl->or_lo = 0; ctl->and_lo = ~0ull; ctl->or_hi = 0; ctl->and_hi = ~0u;
    ctl->part_min = ~0u; ctl->part_max = 0;
    ctl->npasses = 0;
    for (int i = 0; i < 56; ++i) (&ctl->occ[0][0])[i] = 0;
    ctl->n_tie = 0; ctl->tie_overflow = 0; ctl->tie_small = 0; ctl->rows_ticket = 0;
    ctl->n_out = 0; ctl->key_bytes = 0;
    ctl->max_tie_len = 0;
    for (int i = 0; i < MAX_PASSES; ++i) ctl->tickets[i] = 0;
}

constexpr uint32_t MAX_WORLD = 1024;

// shuffle: count records / long-key bytes per destination rank (per-CTA tallies in shared
// memory, then one atomic per destination per CTA)
__global__ void __launch_bounds__(256) export_count_kernel(EntryArrays ent, uint32_t world,
                                                           unsigned long long *nrec,
                                                           unsigned long long *nbytes, const Ctl *ctl)
{
    __shared__ unsigned int s_rec[MAX_WORLD];
    __shared__ unsigned long long s_bytes[MAX_WORLD];
    for (uint32_t d = threadIdx.x; d < world; d += blockDim.x) { s_rec[d] = 0; s_bytes[d] = 0; }
    __syncthreads();
    const uint64_t n = ctl->n;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t d = ent.owner(i);
        atomicAdd(&s_rec[d], 1u);
        Key k = ent.key(i);
        if (key_is_long(k)) atomicAdd(&s_bytes[d], (unsigned long long)key_len(k));
    }
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < world; d += blockDim.x) {
        if (s_rec[d]) atomicAdd(&nrec[d], (unsigned long long)s_rec[d]);
        if (s_bytes[d]) atomicAdd(&nbytes[d], s_bytes[d]);
    }
}

// segment layout per destination: [records 24 B each {k0, k1, count}][long-key bytes]; long
// keys are rewritten to offsets relative to the start of the segment's byte block.  Each CTA
// works on one contiguous chunk of entries: tally per destination, reserve its ranges with
// one atomic per destination, then hand out slots from shared-memory cursors.
__global__ void __launch_bounds__(256) export_scatter_kernel(EntryArrays ent, const uint8_t *bytes,
                                                             uint32_t world, uint8_t *send,
                                                             const unsigned long long *seg_off,
                                                             const unsigned long long *nrec_total,
                                                             unsigned long long *rec_cur,
                                                             unsigned long long *byte_cur,
                                                             const unsigned long long *seg_end, Ctl *ctl)
{
    __shared__ unsigned int s_rec[MAX_WORLD];
    __shared__ unsigned long long s_bytes[MAX_WORLD], s_rec_base[MAX_WORLD], s_byte_base[MAX_WORLD];
    if (ctl->err & ERR_SHUFFLE_FULL) return; // a segment does not fit its slot: reported by the host
    if (seg_end) {
        // exact-size path: the segment bounds came from the HOST's copy of the tallies; if that copy was
        // stale or wrong (read before the count kernel ran), nothing may be scattered -- every CTA
        // sees the same numbers and takes the same way out
        const unsigned long long *nbytes_total = nrec_total + world;
        bool bad = false;
        for (uint32_t d = threadIdx.x; d < world; d += blockDim.x)
            bad |= seg_off[d] + nrec_total[d] * REC_BYTES + nbytes_total[d] > seg_end[d];
        if (__syncthreads_or(bad)) {
            if (threadIdx.x == 0) atomicOr(&ctl->err, ERR_SHUFFLE_FULL);
            return;
        }
    }
    const uint64_t n = ctl->n;
    const uint64_t chunk = (n + gridDim.x - 1) / gridDim.x;
    const uint64_t lo = (uint64_t)blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
    if (lo >= hi) return;
    for (uint32_t d = threadIdx.x; d < world; d += blockDim.x) { s_rec[d] = 0; s_bytes[d] = 0; }
    __syncthreads();
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint32_t d = ent.owner(i);
        atomicAdd(&s_rec[d], 1u);
        Key k = ent.key(i);
        if (key_is_long(k)) atomicAdd(&s_bytes[d], (unsigned long long)key_len(k));
    }
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < world; d += blockDim.x) {
        s_rec_base[d] = s_rec[d] ? atomicAdd(&rec_cur[d], (unsigned long long)s_rec[d]) : 0;
        s_byte_base[d] = s_bytes[d] ? atomicAdd(&byte_cur[d], s_bytes[d]) : 0;
        s_rec[d] = 0; s_bytes[d] = 0;
    }
    __syncthreads();
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint32_t d = ent.owner(i);
        Key k = ent.key(i);
        uint8_t *seg = send + seg_off[d];
        unsigned long long r = s_rec_base[d] + atomicAdd(&s_rec[d], 1u);
        if (key_is_long(k)) {
            uint32_t len = key_len(k);
            unsigned long long bo = s_byte_base[d] + atomicAdd(&s_bytes[d], (unsigned long long)len);
            uint8_t *dst = seg + nrec_total[d] * REC_BYTES + bo;
            uint64_t off = key_off(k);
            for (uint32_t j = 0; j < len; ++j) dst[j] = bytes[off + j];
            k = make_long_key(bo, len, (uint32_t)k.k1);
        }
        ((unsigned long long *)seg)[3 * r] = k.k0;
        ((unsigned long long *)seg)[3 * r + 1] = k.k1;
        ((unsigned long long *)seg)[3 * r + 2] = ent.cnt(i);
    }
}

// fixed-slot shuffle: every destination's segment sits in a slot of slot_bytes, led by a
// 16-byte header {records, long-key bytes}; sizes never visit the host
__global__ void export_header_kernel(uint8_t *send, uint32_t world, unsigned long long slot_bytes,
                                     const unsigned long long *nrec, const unsigned long long *nbytes,
                                     unsigned long long *seg_off, Ctl *ctl)
{
    const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= world) return;
    seg_off[d] = (unsigned long long)d * slot_bytes + 16;
    unsigned long long *hdr = (unsigned long long *)(send + (unsigned long long)d * slot_bytes);
    if (16 + nrec[d] * REC_BYTES + nbytes[d] > slot_bytes) {
        atomicOr(&ctl->err, ERR_SHUFFLE_FULL);
        hdr[0] = ~0ull; hdr[1] = 0; // poison: the receiver reports the overflow too, so every rank fails alike
    } else {
        hdr[0] = nrec[d]; hdr[1] = nbytes[d];
    }
}

// received slots -> entries (all sources in one launch); sets ctl->n
constexpr uint32_t IMPORT_CHUNKS = 32;
__global__ void __launch_bounds__(256) import_fixed_kernel(const uint8_t *recv, uint32_t world,
                                                           unsigned long long slot_bytes, EntryArrays ent,
                                                           unsigned long long max_keys, Ctl *ctl)
{
    __shared__ unsigned long long s_base, s_n, s_total, s_nbytes;
    const uint32_t src = blockIdx.x / IMPORT_CHUNKS, chunk = blockIdx.x % IMPORT_CHUNKS;
    if (threadIdx.x == 0) {
        unsigned long long base = 0, total = 0, mine = 0, mine_bytes = 0;
        for (uint32_t q = 0; q < world; ++q) {
            const unsigned long long *hdr = (const unsigned long long *)(recv + (unsigned long long)q * slot_bytes);
            unsigned long long n = hdr[0], b = hdr[1];
            if (16 + n * REC_BYTES + b > slot_bytes) { n = 0; atomicOr(&ctl->err, ERR_SHUFFLE_FULL); }
            if (q < src) base += n;
            if (q == src) { mine = n; mine_bytes = n ? b : 0; }
            total += n;
        }
        if (total > max_keys) { atomicOr(&ctl->err, ERR_KEYS_FULL); total = max_keys; }
        s_base = base; s_n = mine; s_total = total; s_nbytes = mine_bytes;
        if (blockIdx.x == 0) ctl->n = total;
    }
    __syncthreads();
    const unsigned long long n = s_n, base = s_base;
    const unsigned long long seg = (unsigned long long)src * slot_bytes + 16;
    const unsigned long long *rec = (const unsigned long long *)(recv + seg);
    const unsigned long long bytes_at = seg + n * REC_BYTES;
    const unsigned long long per = (n + IMPORT_CHUNKS - 1) / IMPORT_CHUNKS;
    const unsigned long long lo = per * chunk, hi = lo + per < n ? lo + per : n;
    for (unsigned long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        if (base + i >= s_total) break;
        Key k; k.k0 = rec[3 * i]; k.k1 = rec[3 * i + 1];
        if (key_is_long(k)) {
            // offsets inside a record are the sender's word: they must stay inside its byte block
            uint64_t off = key_off(k);
            uint32_t len = key_len(k);
            if (off + len > s_nbytes) { atomicOr(&ctl->err, ERR_SHUFFLE_FULL); off = 0; len = (uint32_t)min((unsigned long long)len, s_nbytes); }
            k = make_long_key(bytes_at + off, len, (uint32_t)k.k1);
        }
        GEntry ge; ge.k0 = k.k0; ge.k1 = k.k1; ge.cnt = rec[3 * i + 2]; ge.pad = 0;
        ent.e[base + i] = ge;
    }
}

// received segment -> entries; long keys rebased to offsets inside the receive buffer
__global__ void __launch_bounds__(256) import_segment_kernel(const uint8_t *recv, uint64_t seg_off,
                                                             uint64_t nrec, uint64_t nbytes, uint64_t ent_base,
                                                             EntryArrays ent, Ctl *ctl)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const unsigned long long *rec = (const unsigned long long *)(recv + seg_off);
    const uint64_t bytes_at = seg_off + nrec * REC_BYTES;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nrec; i += stride) {
        Key k; k.k0 = rec[3 * i]; k.k1 = rec[3 * i + 1];
        if (key_is_long(k)) {
            uint64_t off = key_off(k);
            uint32_t len = key_len(k);
            if (off + len > nbytes) { atomicOr(&ctl->err, ERR_SHUFFLE_FULL); off = 0; len = (uint32_t)min((unsigned long long)len, (unsigned long long)nbytes); }
            k = make_long_key(bytes_at + off, len, (uint32_t)k.k1);
        }
        GEntry ge; ge.k0 = k.k0; ge.k1 = k.k1; ge.cnt = rec[3 * i + 2]; ge.pad = 0;
        ent.e[ent_base + i] = ge;
    }
}

} // namespace

struct mr_result {
    mr_ctx *ctx;
    uint64_t first, n, cur;
    std::vector<uint64_t> count, key_off;
    std::vector<uint8_t> key_bytes;
};

struct mr_ctx {
    mr_config cfg{};
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    int sm_count = 148;
    int map_variant = 0;

    // input arena
    uint8_t *d_arena = nullptr;
    uint64_t arena_cap = 0;   // usable bytes (allocation has ARENA_SLACK more)
    uint64_t used = 0;        // bytes staged so far (end of the last split)
    uint64_t mapped_segs = 0; // segments already consumed by mr_map_finish
    uint64_t input_bytes = 0;
    bool any_split = false;
    uint8_t *d_pool = nullptr; // bigram contexts: bytes of keys longer than 15 bytes
    uint8_t *h_file = nullptr; // pinned: mr_map_split_path reads files through it
    uint64_t file_cap = 0;
    bool file_busy = false;   // a copy out of h_file may still be in flight
    Gap *h_gaps = nullptr;    // pinned
    Gap *d_gaps = nullptr;
    uint32_t ngaps = 0;       // gaps collected since the last flush, at h_gaps[gap_base ..]
    uint32_t gap_base = 0;    // the pinned list is a ring: earlier flushes may still be reading below this

    // tables and entries
    GEntry *d_gtab = nullptr;
    bool table_dirty = true;  // the table may hold keys (set by the map kernels, cleared by compaction / sender)
    uint64_t slots = 0;
    uint64_t max_keys = 0;
    EntryArrays ent{};
    SortBufs bufs{};
    uint32_t *d_status = nullptr;
    uint32_t max_tiles = 0;
    TieItem *d_ties = nullptr;
    uint32_t *d_posv = nullptr;
    RowsStatus rows_st{};
    uint32_t rows_tiles = 0;
    RowArrays rows{};
    uint64_t key_bytes_cap = 0;
    Ctl *d_ctl = nullptr;
    Ctl *h_ctl = nullptr; // pinned
    MsdBins *d_msd = nullptr; // [2], alternating by job (msd_kernels.cuh)
    uint32_t msd_parity = 0;
    bool prefer_lsd = false;  // MRGPU_REDUCE=lsd: always the LSD path
    unsigned reader_configured = 0; // bit 0 / 1: shared-memory carveout of the compaction / p2p sender set on this device
    uint32_t lsd_jobs_left = 0; // non-zero: the MSD path declined a job of this context; its jobs take the LSD path at once until the owners change
    bool used_msd = false;    // this job's reduce side was launched on the MSD path
    unsigned long long *h_stage = nullptr; // pinned, [MAX_WORLD + 1]
    const uint8_t *bytes_base = nullptr;

    // shuffle
    uint8_t *d_send = nullptr;
    uint64_t send_cap = 0;
    uint8_t *d_recv = nullptr;
    uint64_t recv_cap = 0;
    unsigned long long *d_shuf = nullptr; // [4][world]: nrec, nbytes, rec_cur, byte_cur  + [world+1] seg_off
    std::vector<uint64_t> send_off, send_nrec;
    bool imported = false;
    bool shuffle_counted = false;
    void *nccl_comm = nullptr; // mr_nccl_init: communicator of the library's own transport

    // peer-to-peer shuffle (mr_p2p_setup / mr_p2p_attach / mr_shuffle_p2p), layout in p2p_kernels.cuh
    GEntry *ent_own = nullptr;          // the context's own entry array (ent.e points into the window after a p2p shuffle)
    uint8_t *p2p_win = nullptr;         // this rank's window
    uint64_t p2p_slot = 0, p2p_win_bytes = 0;
    std::vector<uint8_t *> p2p_peer;    // every rank's window as mapped in this process
    std::vector<char> p2p_ipc;          // 1: opened with cudaIpcOpenMemHandle (to be closed)
    uint8_t **d_p2p_buf = nullptr;      // device [2][world]: the peers' buffers per parity
    uint8_t **d_p2p_win = nullptr;      // device [world]
    unsigned int *d_p2p_done = nullptr;
    uint64_t p2p_epoch = 0;
    bool p2p_attached = false;
    bool p2p_sent_pending = false;      // Ctl::sent holds the records this rank sent (map_keys)

    // owner ranks by key range (mr_shuffle_set_owners); none: owner = partition mod world_size
    unsigned long long *d_split_word = nullptr;
    uint32_t *d_split_part = nullptr;
    std::vector<uint64_t> split_word;
    std::vector<uint32_t> split_part;
    OwnerMap owners() const
    {
        OwnerMap m;
        m.word = d_split_word; m.part = d_split_part; m.n = (uint32_t)split_word.size(); m.world = cfg.world_size;
        return m;
    }

    // state
    enum { IDLE, MAPPED, REDUCED, SYNCED } state = IDLE;
    bool map_dirty = false; // splits staged since the last mr_map_finish
    bool has_pairs = false; // the arena holds mr_map_pairs blobs (not text: the splits cannot be re-mapped)
    mr_job_info info{};
    uint32_t tie_rounds = 0;
    cudaEvent_t ev_map0 = nullptr, ev_map1 = nullptr, ev_red0 = nullptr, ev_red1 = nullptr, ev_gaps = nullptr;
    bool gaps_pending = false;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy = nullptr, ev_job = nullptr;
    bool copies_pending = false, copy_synced = false;
    uint64_t staged_since_map = 0, auto_map_bytes = 32ull << 20;
    uint64_t remap_upto = 0;  // keep_input: segments that hold the resident splits
    struct RecSplit { uint64_t off_k, off_v, n; };
    std::vector<RecSplit> rec_splits; // MR_MAP_RECORDS_U32: staged record batches
    bool rec_remap = false;
    bool map_timed = false, red_timed = false;
    mr_stats stats{};

    // host copy of the rows for cursors
    bool rows_cached = false;
    std::vector<uint32_t> h_part;
    std::vector<uint64_t> h_count, h_key_off, h_part_begin;
    std::vector<uint8_t> h_key_bytes;
};

namespace {

int fail(mr_ctx *c, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}

#define CU(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(ctx, MR_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                                      \
    } while (0)

template <typename T>
cudaError_t dmalloc(T **p, uint64_t count)
{
    return cudaMalloc((void **)p, std::max<uint64_t>(count, 1) * sizeof(T));
}

inline unsigned grid_for(uint64_t n, unsigned threads, unsigned cap)
{
    uint64_t g = (n + threads - 1) / threads;
    return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(g, cap));
}

template <int WARPS, int LOG2>
cudaError_t launch_map(mr_ctx *ctx, const MapParams &p)
{
    size_t smem = MapSmemLayout<LOG2>::total(WARPS);
    cudaError_t ea = cudaFuncSetAttribute(map_combine_kernel<WARPS, LOG2>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (ea != cudaSuccess) return ea;
    if (const char *cv = getenv("MRGPU_MAP_CARVEOUT"))   // measurement switch: % of the SM's 228 KB kept as shared memory
        cudaFuncSetAttribute(map_combine_kernel<WARPS, LOG2>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(cv));
    int ctas_per_sm = (int)(227 * 1024 / (smem + 1024));
    ctas_per_sm = std::max(1, std::min(ctas_per_sm, 2048 / (WARPS * 32)));
    uint64_t nwarps_needed = p.nseg;
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count * ctas_per_sm,
                                                 (nwarps_needed + WARPS - 1) / WARPS);
    grid = std::max(grid, 1u);
    map_combine_kernel<WARPS, LOG2><<<grid, WARPS * 32, smem, ctx->stream>>>(p);
    ctx->stats.launches++;
    return cudaGetLastError();
}

int run_general_ties(mr_ctx *ctx, uint32_t nt);
int run_lsd_ties(mr_ctx *ctx, uint32_t max_len);
int launch_rows(mr_ctx *ctx, bool zeroed);
void nccl_release(mr_ctx *ctx);
void p2p_release(mr_ctx *ctx);

} // namespace

// All entry points are declared extern "C" in mrgpu.h.

int mr_abi_version(void) { return MRGPU_ABI_VERSION; }

const char *mr_last_error(const mr_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

uint32_t mr_partition_of(const void *key, uint64_t klen, uint32_t num_reducers)
{
    // host restatement of FnvLua52 (common.cuh), same IEEE operations
    const uint8_t *k = (const uint8_t *)key;
    volatile double h = 2166136261.0;
    for (uint64_t i = 0; i < klen; ++i) {
        volatile double p = h * 16777619.0;
        volatile double q = std::floor(p * (1.0 / 4294967296.0));
        double m = p - q * 4294967296.0;
        h = (double)((uint32_t)m ^ (uint32_t)k[i]);
    }
    double r = (double)num_reducers, hh = h;
    volatile double q = std::floor(hh / r);
    return (uint32_t)(hh - q * r);
}

void mr_ctx_destroy(mr_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    nccl_release(ctx);
    p2p_release(ctx);
    cudaFree(ctx->d_split_word); cudaFree(ctx->d_split_part);
    cudaFree(ctx->d_arena); cudaFree(ctx->d_gtab); cudaFree(ctx->d_pool);
    cudaFree(ctx->ent_own);
    for (int i = 0; i < 2; ++i) cudaFree(ctx->bufs.item[i]);
    cudaFree(ctx->d_status); cudaFree(ctx->d_ties); cudaFree(ctx->d_posv);
    cudaFree(ctx->rows_st.heads);
    cudaFree(ctx->rows.part); cudaFree(ctx->rows.count); cudaFree(ctx->rows.key_off);
    cudaFree(ctx->rows.key_bytes);
    cudaFree(ctx->rows.part_begin);
    cudaFree(ctx->d_msd);
    cudaFree(ctx->d_ctl); cudaFree(ctx->d_send); cudaFree(ctx->d_recv); cudaFree(ctx->d_shuf);
    if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->h_gaps) cudaFreeHost(ctx->h_gaps);
    if (ctx->h_file) cudaFreeHost(ctx->h_file);
    cudaFree(ctx->d_gaps);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    for (cudaEvent_t e : { ctx->ev_map0, ctx->ev_map1, ctx->ev_red0, ctx->ev_red1, ctx->ev_gaps, ctx->ev_copy, ctx->ev_job })
        if (e) cudaEventDestroy(e);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int mr_ctx_create(const mr_config *cfg, mr_ctx **out)
{
    mr_ctx *ctx = nullptr; // CU() reports into g_create_error while ctx is null
    if (!cfg || !out) return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: null argument");
    *out = nullptr;
    if (cfg->struct_size != sizeof(mr_config))
        return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: struct_size %u != %zu", cfg->struct_size,
                    sizeof(mr_config));
    const bool text_job = (cfg->map_kind == MR_MAP_TOKENIZE_WS || cfg->map_kind == MR_MAP_BIGRAM_WS) &&
                          cfg->partition_kind == MR_PARTITION_FNV1_LUA52;
    const bool rec_job = cfg->map_kind == MR_MAP_RECORDS_U32 && cfg->partition_kind == MR_PARTITION_MOD_U32;
    // text jobs: every token is emit(word, 1), so a count IS the sum; a minimum / maximum of ones says
    // nothing, so those reducers take their values from mr_map_pairs only (string keys with values)
    const bool red_ok = cfg->reduce_kind == MR_REDUCE_SUM_U64 || (rec_job && cfg->reduce_kind <= MR_REDUCE_COUNT) ||
                        (text_job && cfg->reduce_kind == MR_REDUCE_COUNT) ||
                        (cfg->map_kind == MR_MAP_TOKENIZE_WS && text_job &&
                         (cfg->reduce_kind == MR_REDUCE_MIN_U64 || cfg->reduce_kind == MR_REDUCE_MAX_U64));
    if (!(text_job || rec_job) || !red_ok)
        return fail(nullptr, MR_ERR_UNSUPPORTED, "mr_ctx_create: unsupported map/reduce/partition kind");
    if (cfg->num_reducers == 0 || cfg->num_reducers > (1u << 24))
        return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: num_reducers must be in 1..2^24");
    if (cfg->world_size == 0 || cfg->rank >= cfg->world_size || cfg->world_size > MAX_WORLD)
        return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: bad world_size/rank");
    if (cfg->arena_bytes == 0 || cfg->arena_bytes + ARENA_SLACK >= MAX_BYTE_SPACE)
        return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: arena_bytes must be in 1..2^37");
    int ndev = 0;
    cudaError_t e0 = cudaGetDeviceCount(&ndev);
    if (e0 != cudaSuccess || ndev == 0)
        return fail(nullptr, MR_ERR_CUDA, "mr_ctx_create: no CUDA device (%s); there is no CPU path",
                    cudaGetErrorString(e0));
    if (cfg->device < 0 || cfg->device >= ndev)
        return fail(nullptr, MR_ERR_INVALID, "mr_ctx_create: device %d out of range", cfg->device);
    CU(cudaSetDevice(cfg->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return fail(nullptr, MR_ERR_UNSUPPORTED,
                    "mr_ctx_create: device is sm_%d%d, this library is built for sm_100a only", prop.major,
                    prop.minor);

    ctx = new mr_ctx();
    ctx->cfg = *cfg;
    ctx->sm_count = prop.multiProcessorCount;
    if (const char *v = getenv("MRGPU_MAP_VARIANT")) ctx->map_variant = atoi(v);
    // room for the caller's bytes plus what each map round pads (a segment and the guard)
    ctx->arena_cap = round_up(cfg->arena_bytes, SEG) + SEG + (cfg->arena_bytes / (16ull << 20) + 8) * (SEG + 2 * TAIL_GUARD);
    uint64_t mk = cfg->max_keys;
    if (mk == 0) mk = std::min<uint64_t>(std::max<uint64_t>(cfg->arena_bytes / 64, 1u << 16), 1u << 24);
    if (mk >= (1ull << 30)) { delete ctx; return fail(nullptr, MR_ERR_INVALID, "max_keys must be < 2^30"); }
    ctx->max_keys = mk;
    uint64_t slots = cfg->table_slots;
    if (slots == 0) { slots = 1; while (slots < 2 * mk) slots <<= 1; }
    if (slots & (slots - 1) || slots > (1ull << 31) || slots < 1024) {
        delete ctx;
        return fail(nullptr, MR_ERR_INVALID, "table_slots must be a power of two in 2^10..2^31");
    }
    ctx->slots = slots;
    ctx->max_tiles = (uint32_t)((mk + SORT_TILE - 1) / SORT_TILE);
    // every input byte is in at most one word-count key, but in up to two bigram keys
    const uint64_t key_space = cfg->map_kind == MR_MAP_BIGRAM_WS ? 2 * ctx->arena_cap : ctx->arena_cap;
    ctx->key_bytes_cap = std::min<uint64_t>(key_space, mk * 64) + 64;

#define CUC(call)                                                                                    \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            std::string m = std::string(#call) + " failed: " + cudaGetErrorString(e_);               \
            mr_ctx_destroy(ctx);                                                                     \
            return fail(nullptr, e_ == cudaErrorMemoryAllocation ? MR_ERR_NOMEM : MR_ERR_CUDA, "%s", \
                        m.c_str());                                                                  \
        }                                                                                            \
    } while (0)

    if (cfg->stream) ctx->stream = (cudaStream_t)cfg->stream;
    else { CUC(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true; }
    CUC(dmalloc(&ctx->d_arena, ctx->arena_cap + ARENA_SLACK));
    CUC(dmalloc(&ctx->d_gtab, slots));
    if (cfg->map_kind == MR_MAP_BIGRAM_WS) CUC(dmalloc(&ctx->d_pool, 2 * ctx->arena_cap + ARENA_SLACK));
    CUC(dmalloc(&ctx->ent.e, mk));
    ctx->ent_own = ctx->ent.e;
    for (int i = 0; i < 2; ++i) {
        CUC(dmalloc(&ctx->bufs.item[i], mk));
    }
    CUC(dmalloc(&ctx->d_status, (uint64_t)MAX_PASSES * ctx->max_tiles * RADIX_BINS));
    uint64_t tie_cap = 1; while (tie_cap < mk) tie_cap <<= 1;
    CUC(dmalloc(&ctx->d_ties, tie_cap)); CUC(dmalloc(&ctx->d_posv, tie_cap));
    ctx->rows_tiles = (uint32_t)((mk + UNIT_W - 1) / UNIT_W) + 2; // look-back words: one per rows tile or per MSD unit
    CUC(dmalloc(&ctx->rows_st.heads, 3ull * ctx->rows_tiles));
    ctx->rows_st.bytes_agg = ctx->rows_st.heads + ctx->rows_tiles;
    ctx->rows_st.bytes_inc = ctx->rows_st.heads + 2ull * ctx->rows_tiles;
    CUC(dmalloc(&ctx->rows.part, mk)); CUC(dmalloc(&ctx->rows.count, mk));
    CUC(dmalloc(&ctx->rows.key_off, mk + 1)); CUC(dmalloc(&ctx->rows.key_bytes, ctx->key_bytes_cap));
    CUC(dmalloc(&ctx->rows.part_begin, (uint64_t)cfg->num_reducers + 1));
    CUC(dmalloc(&ctx->d_ctl, 1));
    CUC(dmalloc(&ctx->d_msd, 2));
    // on the context's stream: a cudaMemset on the default stream is not ordered against a non-blocking
    // stream, and the first job's histogram kernel counts into this block
    CUC(cudaMemsetAsync(ctx->d_msd, 0, 2 * sizeof(MsdBins), ctx->stream));
    if (const char *v = getenv("MRGPU_REDUCE")) ctx->prefer_lsd = strcmp(v, "lsd") == 0;
    CUC(cudaMallocHost((void **)&ctx->h_ctl, sizeof(Ctl)));
    CUC(cudaMallocHost((void **)&ctx->h_stage, (MAX_WORLD + 1) * 8));
    CUC(cudaMallocHost((void **)&ctx->h_gaps, (MAX_GAPS + 1) * sizeof(Gap)));
    CUC(dmalloc(&ctx->d_gaps, MAX_GAPS + 1));
    CUC(dmalloc(&ctx->d_shuf, 5ull * cfg->world_size + 2)); // last word: largest segment (mr_shuffle_nccl)
    CUC(dmalloc(&ctx->d_split_word, cfg->world_size));
    CUC(dmalloc(&ctx->d_split_part, cfg->world_size));
    CUC(cudaEventCreate(&ctx->ev_map0)); CUC(cudaEventCreate(&ctx->ev_map1));
    CUC(cudaEventCreate(&ctx->ev_red0)); CUC(cudaEventCreate(&ctx->ev_red1));
    CUC(cudaEventCreateWithFlags(&ctx->ev_gaps, cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&ctx->ev_job, cudaEventDisableTiming));
    CUC(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    if (const char *v = getenv("MRGPU_AUTO_MAP_MB")) ctx->auto_map_bytes = (uint64_t)atoll(v) << 20;
    ctx->bytes_base = ctx->d_pool ? ctx->d_pool : ctx->d_arena;
    int rc = mr_job_begin(ctx, 0);
    if (rc != MR_OK) { std::string m = ctx->err; mr_ctx_destroy(ctx); return fail(nullptr, rc, "%s", m.c_str()); }
    *out = ctx;
    return MR_OK;
}

int mr_job_begin(mr_ctx *ctx, int keep_input)
{
    if (!ctx) return MR_ERR_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (keep_input && ctx->has_pairs)
        return fail(ctx, MR_ERR_UNSUPPORTED, "mr_job_begin: keep_input is not available after mr_map_pairs (the ingested runs are not kept as text)");
    if (!keep_input) {
        ctx->has_pairs = false;
        ctx->used = 0; ctx->mapped_segs = 0; ctx->input_bytes = 0; ctx->any_split = false;
        ctx->map_dirty = false;
        ctx->ngaps = 0;
        ctx->copy_synced = false;
        ctx->staged_since_map = 0;
        ctx->remap_upto = 0;
        ctx->rec_splits.clear();
        ctx->rec_remap = false;
    } else {
        // keep the staged splits: the next mr_map_finish maps all of them again
        ctx->remap_upto = std::max(ctx->remap_upto, ctx->mapped_segs);
        ctx->mapped_segs = 0;
        ctx->map_dirty = ctx->any_split && ctx->cfg.map_kind != MR_MAP_RECORDS_U32;
        ctx->rec_remap = !ctx->rec_splits.empty();
        ctx->staged_since_map = 0;
    }
    // the kernels that read the table out (compaction, peer-to-peer sender) leave it empty; it is
    // cleared here only when a job mapped something and never got that far
    if (ctx->table_dirty) {
        CU(cudaMemsetAsync(ctx->d_gtab, 0, ctx->slots * sizeof(GEntry), ctx->stream));
        ctx->table_dirty = false;
    }
    ctl_job_reset_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->d_ctl);
    CU(cudaGetLastError());
    ctx->bytes_base = ctx->d_pool ? ctx->d_pool : ctx->d_arena;
    ctx->ent.e = ctx->ent_own;
    ctx->imported = false;
    ctx->shuffle_counted = false;
    ctx->state = mr_ctx::IDLE;
    ctx->rows_cached = false;
    ctx->tie_rounds = 0;
    ctx->used_msd = false;
    ctx->map_timed = ctx->red_timed = false;
    ctx->stats = mr_stats{};
    ctx->stats.launches = 1;
    memset(&ctx->info, 0, sizeof ctx->info);
    return MR_OK;
}

static int run_map_round(mr_ctx *ctx);

static int flush_gaps(mr_ctx *ctx)
{
    if (!ctx->ngaps) return MR_OK;
    CU(cudaMemcpyAsync(ctx->d_gaps, ctx->h_gaps + ctx->gap_base, ctx->ngaps * sizeof(Gap), cudaMemcpyHostToDevice, ctx->stream));
    // between splits: a space for the word count; a newline for bigrams (pairs never span splits)
    const uint8_t fill = ctx->cfg.map_kind == MR_MAP_BIGRAM_WS ? 0x0A : 0x20;
    fill_gaps_kernel<<<std::min<uint32_t>(ctx->ngaps, 1024u), 256, 0, ctx->stream>>>(ctx->d_arena, ctx->d_gaps, ctx->ngaps, fill);
    CU(cudaGetLastError());
    CU(cudaEventRecord(ctx->ev_gaps, ctx->stream)); // marks the last read of the pinned list so far
    ctx->gaps_pending = true;
    ctx->stats.launches++;
    ctx->gap_base += ctx->ngaps;
    ctx->ngaps = 0;
    return MR_OK;
}

// Appends to the pinned gap list.  The host never waits for a flush in the steady state (a wait
// here would drain the copy queue once per map round); only when the ring is used up does it
// wait for the flushes behind it and start over at the front.
static int gap_push(mr_ctx *ctx, uint64_t off, uint64_t len)
{
    if (ctx->gap_base + ctx->ngaps == MAX_GAPS) {
        int rc = flush_gaps(ctx);
        if (rc) return rc;
        if (ctx->gaps_pending) { CU(cudaEventSynchronize(ctx->ev_gaps)); ctx->gaps_pending = false; }
        ctx->gap_base = 0;
    }
    Gap &g = ctx->h_gaps[ctx->gap_base + ctx->ngaps++];
    g.off = off;
    g.len = len;
    return MR_OK;
}

static int stage_split(mr_ctx *ctx, const void *bytes, uint64_t n, cudaMemcpyKind kind)
{
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->cfg.map_kind == MR_MAP_RECORDS_U32)
        return fail(ctx, MR_ERR_INVALID, "mr_map_split: this context maps records, use mr_map_records");
    if (ctx->cfg.reduce_kind == MR_REDUCE_MIN_U64 || ctx->cfg.reduce_kind == MR_REDUCE_MAX_U64)
        return fail(ctx, MR_ERR_UNSUPPORTED, "mr_map_split: a tokenised split emits (word, 1); a min/max context takes (key, value) pairs (mr_map_pairs)");
    if (n && !bytes) return fail(ctx, MR_ERR_INVALID, "mr_map_split: null bytes");
    if (ctx->state != mr_ctx::IDLE && ctx->state != mr_ctx::MAPPED)
        return fail(ctx, MR_ERR_INVALID, "mr_map_split: job already reduced; call mr_job_begin");
    CU(cudaSetDevice(ctx->cfg.device));
    // every split starts on a 16-byte boundary, at least one 0x20 after the previous one
    uint64_t off = ctx->any_split ? round_up(ctx->used + 1, 16) : 0;
    if (off + n + SEG > ctx->arena_cap)
        return fail(ctx, MR_ERR_NOMEM, "mr_map_split: arena full (%llu + %llu > %llu bytes)",
                    (unsigned long long)off, (unsigned long long)n, (unsigned long long)ctx->arena_cap);
    if (off > ctx->used) { int rc = gap_push(ctx, ctx->used, off - ctx->used); if (rc) return rc; }
    if (kind == cudaMemcpyHostToDevice) {
        // host->device copies run on their own stream so that they overlap the map rounds
        if (!ctx->copy_synced) {
            // the arena may still be read by the previous job: wait for everything queued so far
            CU(cudaEventRecord(ctx->ev_job, ctx->stream));
            CU(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_job, 0));
            ctx->copy_synced = true;
        }
        if (n) CU(cudaMemcpyAsync(ctx->d_arena + off, bytes, n, kind, ctx->copy_stream));
        ctx->copies_pending = true;
        ctx->stats.h2d_bytes += n;
    } else if (n) {
        CU(cudaMemcpyAsync(ctx->d_arena + off, bytes, n, kind, ctx->stream));
    }
    ctx->used = off + n;
    ctx->input_bytes += n;
    ctx->any_split = true;
    ctx->map_dirty = true;
    ctx->staged_since_map += n;
    if (ctx->auto_map_bytes && ctx->staged_since_map >= ctx->auto_map_bytes) return run_map_round(ctx);
    return MR_OK;
}

int mr_map_split(mr_ctx *ctx, uint64_t split_id, const void *bytes, uint64_t n)
{
    (void)split_id;
    return stage_split(ctx, bytes, n, cudaMemcpyHostToDevice);
}

int mr_map_split_path(mr_ctx *ctx, uint64_t split_id, const char *path)
{
    // what the reference's mapfn does with its value (mapfn.lua:4: io.lines(value)): the file is
    // read here, into a page-locked buffer the context keeps, and staged like any other split
    if (!ctx || !path) return MR_ERR_INVALID;
    FILE *f = fopen(path, "rb");
    if (!f) return fail(ctx, MR_ERR_INVALID, "mr_map_split_path: cannot open %s", path);
    int rc = MR_OK;
    uint64_t n = 0;
    if (fseek(f, 0, SEEK_END) == 0) { long sz = ftell(f); n = sz > 0 ? (uint64_t)sz : 0; }
    rewind(f);
    CU(cudaSetDevice(ctx->cfg.device));
    // the buffer is reused (or replaced): the previous file's copy must have left it
    if (ctx->file_busy) { CU(cudaStreamSynchronize(ctx->copy_stream)); ctx->file_busy = false; }
    if (n > ctx->file_cap) {
        if (ctx->h_file) cudaFreeHost(ctx->h_file);
        ctx->h_file = nullptr; ctx->file_cap = 0;
        cudaError_t e = cudaMallocHost((void **)&ctx->h_file, n + n / 4 + 4096);
        if (e != cudaSuccess) { fclose(f); return fail(ctx, MR_ERR_NOMEM, "mr_map_split_path: %s", cudaGetErrorString(e)); }
        ctx->file_cap = n + n / 4 + 4096;
    }
    if (n && fread(ctx->h_file, 1, n, f) != n) rc = fail(ctx, MR_ERR_INVALID, "mr_map_split_path: short read of %s", path);
    fclose(f);
    if (rc) return rc;
    (void)split_id;
    ctx->file_busy = n != 0;
    return stage_split(ctx, ctx->h_file, n, cudaMemcpyHostToDevice);
}

int mr_map_split_device(mr_ctx *ctx, uint64_t split_id, const void *dev_bytes, uint64_t n)
{
    (void)split_id;
    return stage_split(ctx, dev_bytes, n, cudaMemcpyDeviceToDevice);
}

// Maps every segment staged since the last round (tokenise + combine into the global table).
static int run_map_round(mr_ctx *ctx)
{
    if (!ctx->map_dirty) return MR_OK;
    if (ctx->copies_pending) {
        CU(cudaEventRecord(ctx->ev_copy, ctx->copy_stream));
        CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_copy, 0));
        ctx->copies_pending = false;
    }
    const bool remap_only = ctx->remap_upto && ctx->staged_since_map == 0 && ctx->ngaps == 0;
    uint64_t end;
    if (remap_only) {
        end = ctx->remap_upto * SEG; // resident splits, already padded: map exactly what was mapped before
    } else {
        // pad the tail with spaces up to a whole segment plus the look-ahead window
        end = round_up(std::max<uint64_t>(ctx->used, 1), SEG);
        { int rc = gap_push(ctx, ctx->used, end + TAIL_GUARD - ctx->used); if (rc) return rc; }
        { int rc = flush_gaps(ctx); if (rc) return rc; }
    }
    if (ctx->cfg.map_kind == MR_MAP_BIGRAM_WS) {
        BigramParams bp;
        bp.arena = ctx->d_arena;
        bp.begin = ctx->mapped_segs * SEG;
        bp.end = end;
        bp.limit = end + TAIL_GUARD;
        bp.gtab = ctx->d_gtab;
        ctx->table_dirty = true;
        bp.gmask = (uint32_t)(ctx->slots - 1);
        bp.max_key = MR_MAX_TOKEN_BYTES;
        bp.pool = ctx->d_pool;
        bp.pool_used = &ctx->d_ctl->pool_used;
        bp.pool_cap = 2 * ctx->arena_cap;
        bp.pairs = &ctx->d_ctl->tokens;
        bp.err = &ctx->d_ctl->err;
        if (bp.end > bp.begin) {
            map_bigram_kernel<<<grid_for(bp.end - bp.begin, 256, ctx->sm_count * 8), 256, 0, ctx->stream>>>(bp);
            ctx->stats.launches++;
            CU(cudaGetLastError());
        }
        ctx->mapped_segs = end / SEG;
        if (!remap_only) ctx->used = end + TAIL_GUARD;
        ctx->map_dirty = false;
        ctx->staged_since_map = 0;
        return MR_OK;
    }
    MapParams p;
    p.arena = ctx->d_arena;
    p.first_seg = ctx->mapped_segs;
    p.nseg = end / SEG - ctx->mapped_segs;
    p.gtab = ctx->d_gtab;
    ctx->table_dirty = true;
    p.gmask = (uint32_t)(ctx->slots - 1);
    p.max_token = MR_MAX_TOKEN_BYTES;
    p.tokens = &ctx->d_ctl->tokens;
    p.err = &ctx->d_ctl->err;
    p.dbg = ctx->d_ctl->dbg;
    if (p.nseg) {
        cudaError_t e;
        switch (ctx->map_variant) {
        case 1: e = launch_map<16, 12>(ctx, p); break;
        case 2: e = launch_map<16, 13>(ctx, p); break;
        case 3: e = launch_map<32, 12>(ctx, p); break;
#ifdef MRGPU_V_WARPS
        default: e = launch_map<MRGPU_V_WARPS, MRGPU_V_LOG2>(ctx, p); break;
#else
        default: e = launch_map<32, 13>(ctx, p); break;
#endif
        }
        if (e != cudaSuccess) return fail(ctx, MR_ERR_CUDA, "map kernel launch: %s", cudaGetErrorString(e));
    }
    ctx->mapped_segs = end / SEG;
    // later splits start in fresh segments, past the guard this round's fill kernel writes
    if (!remap_only) ctx->used = end + TAIL_GUARD;
    ctx->map_dirty = false;
    ctx->staged_since_map = 0;
    return MR_OK;
}

static int launch_records(mr_ctx *ctx, const mr_ctx::RecSplit &r)
{
    if (!r.n) return MR_OK;
    RecordsParams p;
    p.keys = (const uint32_t *)(ctx->d_arena + r.off_k);
    p.vals = (const uint32_t *)(ctx->d_arena + r.off_v);
    p.n = r.n;
    p.gtab = ctx->d_gtab;
    ctx->table_dirty = true;
    p.gmask = (uint32_t)(ctx->slots - 1);
    p.records = &ctx->d_ctl->tokens;
    p.err = &ctx->d_ctl->err;
    p.op = ctx->cfg.reduce_kind;
    constexpr int LOG2 = 13;
    const size_t smem = (size_t)(1u << LOG2) * 16;
    CU(cudaFuncSetAttribute(map_records_kernel<LOG2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count, (r.n + 4095) / 4096);
    map_records_kernel<LOG2><<<std::max(grid, 1u), 1024, smem, ctx->stream>>>(p);
    ctx->stats.launches++;
    CU(cudaGetLastError());
    return MR_OK;
}

static int stage_records(mr_ctx *ctx, const uint32_t *keys, const uint32_t *values, uint64_t n, cudaMemcpyKind kind)
{
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->cfg.map_kind != MR_MAP_RECORDS_U32)
        return fail(ctx, MR_ERR_INVALID, "mr_map_records: this context maps text splits, use mr_map_split");
    if (n && (!keys || !values)) return fail(ctx, MR_ERR_INVALID, "mr_map_records: null pointer");
    if (ctx->state != mr_ctx::IDLE && ctx->state != mr_ctx::MAPPED)
        return fail(ctx, MR_ERR_INVALID, "mr_map_records: job already reduced; call mr_job_begin");
    CU(cudaSetDevice(ctx->cfg.device));
    mr_ctx::RecSplit r;
    r.off_k = round_up(ctx->used, 16);
    r.off_v = r.off_k + round_up(4 * n, 16);
    r.n = n;
    const uint64_t end = r.off_v + round_up(4 * n, 16);
    if (end + SEG > ctx->arena_cap)
        return fail(ctx, MR_ERR_NOMEM, "mr_map_records: arena full (%llu > %llu bytes)", (unsigned long long)end,
                    (unsigned long long)ctx->arena_cap);
    cudaStream_t cs = ctx->stream;
    if (kind == cudaMemcpyHostToDevice) {
        if (!ctx->copy_synced) {
            CU(cudaEventRecord(ctx->ev_job, ctx->stream));
            CU(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_job, 0));
            ctx->copy_synced = true;
        }
        cs = ctx->copy_stream;
        ctx->stats.h2d_bytes += 8 * n;
    }
    if (n) {
        CU(cudaMemcpyAsync(ctx->d_arena + r.off_k, keys, 4 * n, kind, cs));
        CU(cudaMemcpyAsync(ctx->d_arena + r.off_v, values, 4 * n, kind, cs));
    }
    if (kind == cudaMemcpyHostToDevice) {
        CU(cudaEventRecord(ctx->ev_copy, ctx->copy_stream));
        CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_copy, 0));
    }
    ctx->used = end;
    ctx->input_bytes += 8 * n;
    ctx->any_split = true;
    ctx->rec_splits.push_back(r);
    return launch_records(ctx, r); // map + combine right away, overlapping the next batch's copy
}

int mr_map_records(mr_ctx *ctx, uint64_t split_id, const uint32_t *keys, const uint32_t *values, uint64_t n)
{
    (void)split_id;
    return stage_records(ctx, keys, values, n, cudaMemcpyHostToDevice);
}

int mr_map_records_device(mr_ctx *ctx, uint64_t split_id, const uint32_t *dev_keys, const uint32_t *dev_values,
                          uint64_t n)
{
    (void)split_id;
    return stage_records(ctx, dev_keys, dev_values, n, cudaMemcpyDeviceToDevice);
}

int mr_map_pairs(mr_ctx *ctx, uint64_t split_id, const void *key_bytes, const uint64_t *key_off, const uint64_t *values,
                 uint64_t n)
{
    (void)split_id;
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->cfg.map_kind != MR_MAP_TOKENIZE_WS)
        return fail(ctx, MR_ERR_UNSUPPORTED, "mr_map_pairs: string-key pairs go with MR_MAP_TOKENIZE_WS contexts (number keys: mr_map_records)");
    if (ctx->cfg.reduce_kind == MR_REDUCE_COUNT)
        return fail(ctx, MR_ERR_UNSUPPORTED, "mr_map_pairs: a count is not its own combiner; ingest counted runs with MR_REDUCE_SUM_U64");
    if (n && (!key_off || !values || (key_off[n] && !key_bytes))) return fail(ctx, MR_ERR_INVALID, "mr_map_pairs: null pointer");
    if (ctx->state != mr_ctx::IDLE && ctx->state != mr_ctx::MAPPED)
        return fail(ctx, MR_ERR_INVALID, "mr_map_pairs: job already reduced; call mr_job_begin");
    if (!n) return MR_OK;
    if (key_off[0] != 0) return fail(ctx, MR_ERR_INVALID, "mr_map_pairs: key_off[0] must be 0");
    CU(cudaSetDevice(ctx->cfg.device));
    // text staged so far is tokenised first: the blob below must never be read as text
    { int rc = run_map_round(ctx); if (rc) return rc; }
    const uint64_t nbytes = key_off[n];
    const uint64_t at = round_up(std::max<uint64_t>(ctx->used, ctx->mapped_segs * SEG), 16);
    const uint64_t off_at = at + round_up(nbytes + 16, 16), val_at = off_at + (n + 1) * 8, blob_end = val_at + n * 8;
    // the blob ends inside a segment whose last byte is a space, so that the tokeniser, when it
    // later starts at the next segment, does not take that segment's first token for a continuation
    uint64_t end = round_up(blob_end + 1, SEG);
    if (end + SEG > ctx->arena_cap)
        return fail(ctx, MR_ERR_NOMEM, "mr_map_pairs: arena full (%llu > %llu bytes)", (unsigned long long)end,
                    (unsigned long long)ctx->arena_cap);
    cudaStream_t s = ctx->stream;
    if (nbytes) CU(cudaMemcpyAsync(ctx->d_arena + at, key_bytes, nbytes, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(ctx->d_arena + off_at, key_off, (n + 1) * 8, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(ctx->d_arena + val_at, values, n * 8, cudaMemcpyHostToDevice, s));
    ctx->stats.h2d_bytes += nbytes + (2 * n + 1) * 8;
    if (at > ctx->used) { int rc = gap_push(ctx, ctx->used, at - ctx->used); if (rc) return rc; }
    { int rc = gap_push(ctx, blob_end, end - blob_end); if (rc) return rc; }
    { int rc = flush_gaps(ctx); if (rc) return rc; }
    PairsParams p;
    p.arena = ctx->d_arena;
    p.bytes_at = at; p.off_at = off_at; p.val_at = val_at; p.n = n;
    p.gtab = ctx->d_gtab;
    p.gmask = (uint32_t)(ctx->slots - 1);
    p.op = ctx->cfg.reduce_kind;
    p.max_key = MR_MAX_TOKEN_BYTES;
    p.pairs = &ctx->d_ctl->tokens;
    p.err = &ctx->d_ctl->err;
    ctx->table_dirty = true;
    map_pairs_kernel<<<grid_for(n, 256, ctx->sm_count * 8), 256, 0, s>>>(p);
    ctx->stats.launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(s)); // the host arrays are only borrowed for the call
    ctx->used = end;
    ctx->mapped_segs = end / SEG;
    ctx->input_bytes += nbytes;
    ctx->any_split = true;
    ctx->has_pairs = true;
    return MR_OK;
}

int mr_map_finish(mr_ctx *ctx)
{
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::IDLE && ctx->state != mr_ctx::MAPPED)
        return fail(ctx, MR_ERR_INVALID, "mr_map_finish: job already reduced; call mr_job_begin");
    CU(cudaSetDevice(ctx->cfg.device));
    if (!ctx->map_timed) CU(cudaEventRecord(ctx->ev_map0, ctx->stream));
    if (ctx->rec_remap) { // keep_input on a records context: map the resident batches again
        for (const auto &r : ctx->rec_splits) { int rc = launch_records(ctx, r); if (rc) return rc; }
        ctx->rec_remap = false;
    }
    { int rc = run_map_round(ctx); if (rc) return rc; }
    CU(cudaEventRecord(ctx->ev_map1, ctx->stream));
    ctx->map_timed = true;
    ctx->state = mr_ctx::MAPPED;
    return MR_OK;
}

namespace {

// The two kernels that read the whole table out (compaction, p2p sender) are streaming reads whose tiles
// sit in ~33-38 KB of shared memory per CTA.  Eight CTAs per SM would put the SM in its 228 KB
// shared-memory configuration, whose 28 KB of L1 cannot hold the loads in flight: 64 MB then stream at
// 2.7 TB/s.  Four persistent CTAs per SM in the 164 KB configuration (92 KB of L1) read it 8 us faster
// (profiles/r02/compact_carveout.txt).  MRGPU_COMPACT_CARVEOUT / MRGPU_COMPACT_PER_SM: measurement switches.
void table_reader_config(mr_ctx *ctx, unsigned which, const void *kernel, unsigned *per_sm)
{
    static const int carveout = getenv("MRGPU_COMPACT_CARVEOUT") ? atoi(getenv("MRGPU_COMPACT_CARVEOUT")) : 72;
    static const unsigned ctas = getenv("MRGPU_COMPACT_PER_SM") ? (unsigned)std::max(1, atoi(getenv("MRGPU_COMPACT_PER_SM"))) : 4u;
    *per_sm = ctas;
    if (ctx->reader_configured & (1u << which)) return;   // the attribute sticks (per kernel and device): once per context
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
    ctx->reader_configured |= 1u << which;
}

// table -> entries with partition ids and owner ranks, sort items, varying-bit masks: one kernel
int launch_compact(mr_ctx *ctx)
{
    const uint64_t tiles = (ctx->slots + 256 * COMPACT_ITEMS - 1) / (256 * COMPACT_ITEMS);
    unsigned per_sm;
    table_reader_config(ctx, 0, (const void *)compact_partition_kernel, &per_sm);
    unsigned grid = (unsigned)std::min<uint64_t>(tiles, (uint64_t)ctx->sm_count * per_sm);
    compact_partition_kernel<<<grid, 256, 0, ctx->stream>>>(
        ctx->d_gtab, ctx->slots, ctx->bytes_base, ctx->cfg.num_reducers, ctx->cfg.partition_kind, ctx->ent,
        ctx->bufs.item[0], ctx->max_keys, ctx->d_ctl, ctx->owners(), (uint32_t *)ctx->d_ties);
    ctx->table_dirty = false;
    ctx->stats.launches += 1;
    CU(cudaGetLastError());
    return MR_OK;
}

// digit histograms + the radix passes over item[0] (the digit plan is made on the device)
int launch_radix(mr_ctx *ctx)
{
    cudaStream_t s = ctx->stream;
    CU(cudaFuncSetAttribute(radix_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem)));
    radix_hist_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 4), 256, 0, s>>>(ctx->bufs.item[0], ctx->max_keys, ctx->d_ctl,
                                                                                    ctx->d_status, ctx->max_tiles);
    ctx->stats.launches += 1;
    // the sort key has 64 chunk bits plus the bits of the largest partition id: more passes than
    // that can never be needed (which ones are needed is decided on the device)
    uint32_t part_bits = 0;
    while (part_bits < 32 && ((uint64_t)1 << part_bits) < ctx->cfg.num_reducers) ++part_bits;
    const uint32_t max_passes = std::min<uint32_t>(MAX_PASSES, (64 + part_bits + RADIX_BITS - 1) / RADIX_BITS);
    for (uint32_t pass = 0; pass < max_passes; ++pass) {
        radix_pass_kernel<<<ctx->max_tiles, SORT_THREADS, sizeof(SortSmem), s>>>(ctx->bufs, pass, ctx->d_status,
                                                                                ctx->max_tiles, ctx->d_ctl);
        ctx->stats.launches++;
    }
    CU(cudaGetLastError());
    return MR_OK;
}

int launch_sort(mr_ctx *ctx)
{
    cudaStream_t s = ctx->stream;
    { int rc = launch_radix(ctx); if (rc) return rc; }
    uint64_t tie_cap = 1; while (tie_cap < ctx->max_keys) tie_cap <<= 1;
    tie_mark_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 8), 256, 0, s>>>(
        ctx->bufs, ctx->d_ties, (uint32_t)tie_cap, ctx->ent, ctx->bytes_base, ctx->d_ctl, ctx->rows.count, ctx->rows_st.heads,
        3u * ctx->rows_tiles);
    tie_fix_small_kernel<<<1, 1024, 0, s>>>(ctx->bufs, ctx->d_ties, ctx->ent, ctx->bytes_base, ctx->d_ctl);
    ctx->stats.launches += 2;
    CU(cudaGetLastError());
    return MR_OK;
}

// The reduce side in three launches (msd_kernels.cuh): histogram of the most significant digit,
// scatter into bins, then sort + merge + rows per unit of bins in shared memory.
int launch_msd(mr_ctx *ctx)
{
    cudaStream_t s = ctx->stream;
    using SmemD = UnitSmem<UNIT_CAP_DISTINCT>;
    using SmemS = UnitSmem<UNIT_CAP_SHUFFLED>;
    static_assert(4 * (sizeof(SmemD) + 1024) <= 196 * 1024, "four units per SM in the 196 KB shared-memory configuration");
    static_assert(3 * (sizeof(SmemS) + 1024) <= 196 * 1024, "three units per SM in the 196 KB shared-memory configuration");
    CU(cudaFuncSetAttribute(msd_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MSD_SCATTER_SMEM));
    // measurement switches: unused shared memory added to a unit's block (fewer units per SM, smaller carveout)
    static const size_t pad_d = getenv("MRGPU_UNIT_PAD_D") ? (size_t)atoi(getenv("MRGPU_UNIT_PAD_D")) : 0;
    static const size_t pad_s = getenv("MRGPU_UNIT_PAD_S") ? (size_t)atoi(getenv("MRGPU_UNIT_PAD_S")) : 0;
    CU(cudaFuncSetAttribute(msd_unit_kernel<UNIT_CAP_DISTINCT, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(SmemD) + pad_d)));
    CU(cudaFuncSetAttribute(msd_unit_kernel<UNIT_CAP_SHUFFLED, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(SmemS) + pad_s)));
    // no shuffle: the items are the table's entries, one per distinct key, so every item becomes a row
    // (the compaction kernel left the key lengths in d_ties, which only the LSD path's tie kernels use)
    const bool distinct = !ctx->imported && getenv("MRGPU_MSD_LOOKBACK") == nullptr;
    const bool small_units = !ctx->imported;   // after a shuffle every key arrives once per rank: larger units
    MsdBins *mb = ctx->d_msd + ctx->msd_parity, *other = ctx->d_msd + (ctx->msd_parity ^ 1);
    ctx->msd_parity ^= 1;
    msd_hist_kernel<<<grid_for(ctx->max_keys, 256 * 4, ctx->sm_count * 4), 256, 0, s>>>(
        ctx->bufs.item[0], ctx->max_keys, ctx->cfg.num_reducers, ctx->d_ctl, mb, other, ctx->d_posv, ctx->rows_st, ctx->rows_tiles);
    msd_scatter_kernel<<<grid_for(ctx->max_keys, 1024 * 4, ctx->sm_count), 1024, MSD_SCATTER_SMEM, s>>>(
        ctx->bufs.item[0], ctx->d_posv, (MsdItem *)ctx->bufs.item[1], ctx->d_ctl, mb, distinct ? (const uint32_t *)ctx->d_ties : nullptr,
        (small_units ? UNIT_CAP_DISTINCT : UNIT_CAP_SHUFFLED) - UNIT_W);
    const unsigned units = (unsigned)((ctx->max_keys + UNIT_W - 1) / UNIT_W) + 1;
    if (small_units)
        msd_unit_kernel<UNIT_CAP_DISTINCT, 4><<<units, UNIT_THREADS, sizeof(SmemD) + pad_d, s>>>(
            (const MsdItem *)ctx->bufs.item[1], ctx->ent, ctx->bytes_base, ctx->rows, ctx->cfg.num_reducers,
            ctx->cfg.reduce_kind, ctx->key_bytes_cap - 64, ctx->rows_st, ctx->d_ctl, mb, distinct);
    else
        msd_unit_kernel<UNIT_CAP_SHUFFLED, 3><<<units, UNIT_THREADS, sizeof(SmemS) + pad_s, s>>>(
            (const MsdItem *)ctx->bufs.item[1], ctx->ent, ctx->bytes_base, ctx->rows, ctx->cfg.num_reducers,
            ctx->cfg.reduce_kind, ctx->key_bytes_cap - 64, ctx->rows_st, ctx->d_ctl, mb, distinct);
    ctx->stats.launches += 3;
    if (ctx->cfg.reduce_kind == MR_REDUCE_MIN_U64) {
        rows_decode_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 4), 256, 0, s>>>(ctx->rows, ctx->cfg.reduce_kind, ctx->d_ctl);
        ctx->stats.launches += 1;
    }
    CU(cudaGetLastError());
    ctx->used_msd = true;
    return MR_OK;
}

// zeroed: tie_mark_kernel has just cleared what the rows kernel needs cleared (the first pass of a
// job); the second pass after a re-sort clears it here
int launch_rows(mr_ctx *ctx, bool zeroed)
{
    cudaStream_t s = ctx->stream;
    if (!zeroed) {
        CU(cudaMemsetAsync(ctx->rows.count, 0, ctx->max_keys * 8, s));
        CU(cudaMemsetAsync(ctx->rows_st.heads, 0, 3ull * ctx->rows_tiles * 8, s));
        CU(cudaMemsetAsync(&ctx->d_ctl->rows_ticket, 0, sizeof(unsigned int), s));
    }
    rows_kernel<<<ctx->rows_tiles, ROWS_THREADS, 0, s>>>(ctx->bufs, ctx->ent, ctx->bytes_base, ctx->rows,
                                                        ctx->cfg.num_reducers, ctx->cfg.reduce_kind,
                                                        ctx->key_bytes_cap - 64, ctx->rows_st, ctx->d_ctl);
    if (ctx->cfg.reduce_kind == MR_REDUCE_MIN_U64) {
        rows_decode_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 4), 256, 0, s>>>(ctx->rows, ctx->cfg.reduce_kind, ctx->d_ctl);
        ctx->stats.launches += 1;
    }
    ctx->stats.launches += 1;
    CU(cudaGetLastError());
    return MR_OK;
}

// More ties than one CTA can hold: bitonic network over the tie list in global memory.
int run_general_ties(mr_ctx *ctx, uint32_t nt)
{
    cudaStream_t s = ctx->stream;
    uint32_t np2 = 1;
    while (np2 < nt) np2 <<= 1;
    unsigned g = (np2 + 255) / 256;
    tie_pad_kernel<<<g, 256, 0, s>>>(ctx->d_ties, ctx->d_posv, nt, np2);
    ctx->stats.launches++;
    for (uint32_t k = 2; k <= np2; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            tie_bitonic_step_kernel<<<g, 256, 0, s>>>(ctx->d_ties, ctx->d_posv, np2, k, j, ctx->ent, ctx->bytes_base);
            ctx->stats.launches++;
        }
    tie_writeback_kernel<<<(nt + 255) / 256, 256, 0, s>>>(ctx->bufs, ctx->d_ties, ctx->d_posv, nt, ctx->d_ctl);
    ctx->stats.launches++;
    CU(cudaGetLastError());
    return MR_OK;
}

// Many keys in large tie groups: the whole order again, least significant 7-byte round first
// (rekey_kernel's comment).  (max_len - 1) / 7 + 1 rounds of the radix sort.
int run_lsd_ties(mr_ctx *ctx, uint32_t max_len)
{
    cudaStream_t s = ctx->stream;
    const int rmax = max_len ? (int)((max_len - 1) / 7) : 0;
    for (int round = rmax; round >= 0; --round) {
        sort_round_reset_kernel<<<4, 256, 0, s>>>(ctx->d_ctl);
        rekey_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 8), 256, 0, s>>>(ctx->bufs, (uint32_t)round, ctx->ent,
                                                                                     ctx->bytes_base, ctx->max_keys, ctx->d_ctl);
        ctx->stats.launches += 2;
        int rc = launch_radix(ctx);
        if (rc) return rc;
    }
    return MR_OK;
}

int check_device_err(mr_ctx *ctx, unsigned err)
{
    if (err & ERR_TOKEN_TOO_LONG)
        return fail(ctx, MR_ERR_OVERFLOW, "a token is longer than MR_MAX_TOKEN_BYTES (%u)", MR_MAX_TOKEN_BYTES);
    if (err & ERR_TABLE_FULL)
        return fail(ctx, MR_ERR_OVERFLOW, "hash table full (%llu slots): raise table_slots / max_keys",
                    (unsigned long long)ctx->slots);
    if (err & ERR_SHUFFLE_FULL)
        return fail(ctx, MR_ERR_OVERFLOW, "a shuffle segment does not fit the fixed slot: use a larger slot_bytes or the exact-size path");
    if (err & ERR_PEER_TIMEOUT)
        return fail(ctx, MR_ERR_CUDA, "peer-to-peer shuffle: a rank did not deliver its records within %llu s",
                    P2P_TIMEOUT_NS / 1000000000ull);
    if (err & ERR_LOOKBACK_TIMEOUT)
        return fail(ctx, MR_ERR_CUDA, "a sort or rows tile waited for its predecessors for seconds without progress (look-back gave up)");
    if (err & ERR_EMPTY_KEY)
        return fail(ctx, MR_ERR_INVALID, "mr_map_pairs: an empty key (keys have 1 or more bytes)");
    if (err & ERR_KEYS_FULL)
        return fail(ctx, MR_ERR_NOMEM, "more distinct keys (or key bytes) than max_keys=%llu allows",
                    (unsigned long long)ctx->max_keys);
    return MR_OK;
}

} // namespace

int mr_reduce_all(mr_ctx *ctx)
{
    if (!ctx) return MR_ERR_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (ctx->state == mr_ctx::IDLE) { int rc = mr_map_finish(ctx); if (rc) return rc; }
    if (ctx->state != mr_ctx::MAPPED) return fail(ctx, MR_ERR_INVALID, "mr_reduce_all: already reduced");
    if (ctx->map_dirty) { int rc = mr_map_finish(ctx); if (rc) return rc; }
    if (ctx->cfg.world_size > 1 && !ctx->imported)
        return fail(ctx, MR_ERR_INVALID, "mr_reduce_all: this context is rank %u of %u; run one of the shuffles "
                    "(mr_shuffle_p2p, mr_shuffle_nccl, mr_shuffle_count/pack/import) after mr_map_finish first",
                    ctx->cfg.rank, ctx->cfg.world_size);
    CU(cudaEventRecord(ctx->ev_red0, ctx->stream));
    int rc;
    if (!ctx->imported) { if ((rc = launch_compact(ctx))) return rc; }
    if (!ctx->prefer_lsd && ctx->lsd_jobs_left == 0) {
        if ((rc = launch_msd(ctx))) return rc;
    } else {
        if ((rc = launch_sort(ctx))) return rc;
        if ((rc = launch_rows(ctx, true))) return rc;
    }
    CU(cudaEventRecord(ctx->ev_red1, ctx->stream));
    ctx->red_timed = true;
    ctx->state = mr_ctx::REDUCED;
    return MR_OK;
}

int mr_job_sync(mr_ctx *ctx, mr_job_info *info)
{
    if (!ctx) return MR_ERR_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (ctx->state == mr_ctx::IDLE || ctx->state == mr_ctx::MAPPED) { int rc = mr_reduce_all(ctx); if (rc) return rc; }
    if (ctx->state == mr_ctx::REDUCED) {
        CU(cudaMemcpyAsync(ctx->h_ctl, ctx->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        int rc = check_device_err(ctx, ctx->h_ctl->err);
        if (rc) return rc;
        uint32_t flags = ctx->used_msd ? 0u : MR_JOB_REDUCED_LSD;
        if (ctx->used_msd && ctx->h_ctl->msd_fallback) {
            flags = MR_JOB_REDUCED_LSD | MR_JOB_MSD_DECLINED | (ctx->h_ctl->msd_fallback << 8);
            // the MSD path declined (msd_kernels.cuh, "fallback"): the whole reduce side again on the LSD
            // path, from the untouched sort items; later jobs of this context start there
            ctx->lsd_jobs_left = ~0u; // a job shape that does not fit comes back: the LSD path from now on (new owners reset this)
            ctx->used_msd = false;
            CU(cudaMemsetAsync(&ctx->d_ctl->n_tie, 0, offsetof(Ctl, n_out) - offsetof(Ctl, n_tie), ctx->stream));
            if ((rc = launch_sort(ctx))) return rc;
            if ((rc = launch_rows(ctx, true))) return rc;
            CU(cudaEventRecord(ctx->ev_red1, ctx->stream));
            CU(cudaMemcpyAsync(ctx->h_ctl, ctx->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            if ((rc = check_device_err(ctx, ctx->h_ctl->err))) return rc;
        }
        if (ctx->h_ctl->tie_overflow) {
            // rare: many keys agree in their first 7 bytes; order them, then redo the rows
            // keys sharing long prefixes.  Up to 64K of them: pairwise compares of whole keys in a
            // bitonic network; more (and keys of bounded length): the radix sort again, round by round
            const bool lsd = ctx->h_ctl->n_tie > (1u << 16) && ctx->h_ctl->max_tie_len <= 7 * 12;
            if ((rc = lsd ? run_lsd_ties(ctx, ctx->h_ctl->max_tie_len) : run_general_ties(ctx, ctx->h_ctl->n_tie))) return rc;
            if ((rc = launch_rows(ctx, false))) return rc;
                    CU(cudaEventRecord(ctx->ev_red1, ctx->stream));
            CU(cudaMemcpyAsync(ctx->h_ctl, ctx->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            if ((rc = check_device_err(ctx, ctx->h_ctl->err))) return rc;
            ctx->tie_rounds = 2;
        } else {
            ctx->tie_rounds = (ctx->h_ctl->n_tie || ctx->h_ctl->tie_small) ? 1 : 0;
        }
        ctx->stats.d2h_bytes += sizeof(Ctl);
        ctx->info.n_keys = ctx->h_ctl->n_out;
        ctx->info.n_tokens = ctx->h_ctl->tokens;
        ctx->info.input_bytes = ctx->input_bytes;
        ctx->info.key_bytes = ctx->h_ctl->key_bytes;
        if (!ctx->imported) ctx->info.map_keys = ctx->h_ctl->n;
        if (ctx->p2p_sent_pending) { ctx->info.map_keys = ctx->h_ctl->sent; ctx->p2p_sent_pending = false; }
        if (ctx->shuffle_counted && !ctx->p2p_sent_pending && ctx->info.map_keys == 0 && ctx->cfg.world_size > 0) {
            // fixed-slot shuffle: the records this rank sent were only ever counted on the device
            const uint32_t W = ctx->cfg.world_size;
            CU(cudaMemcpyAsync(ctx->h_stage, ctx->d_shuf, W * 8ull, cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            uint64_t sent = 0;
            for (uint32_t d = 0; d < W; ++d) sent += ctx->h_stage[d];
            ctx->info.map_keys = sent;
        }
        ctx->info.tie_rounds = ctx->tie_rounds;
        ctx->info.flags = flags;
        if (ctx->map_timed) cudaEventElapsedTime(&ctx->stats.map_ms, ctx->ev_map0, ctx->ev_map1);
        if (ctx->red_timed) cudaEventElapsedTime(&ctx->stats.reduce_ms, ctx->ev_red0, ctx->ev_red1);
        ctx->state = mr_ctx::SYNCED;
    }
    if (info) *info = ctx->info;
    return MR_OK;
}

int mr_device_rows_get(mr_ctx *ctx, mr_device_rows *rows)
{
    if (!ctx || !rows) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::SYNCED) return fail(ctx, MR_ERR_INVALID, "mr_device_rows_get: call mr_job_sync first");
    rows->n = ctx->info.n_keys;
    rows->part = ctx->rows.part;
    rows->count = (const uint64_t *)ctx->rows.count;
    rows->key_off = (const uint64_t *)ctx->rows.key_off;
    rows->key_bytes = ctx->rows.key_bytes;
    rows->part_begin = (const uint64_t *)ctx->rows.part_begin;
    return MR_OK;
}

int mr_rows_fetch(mr_ctx *ctx, uint32_t *part, uint64_t *count, uint64_t *key_off, uint8_t *key_bytes,
                  uint64_t *part_begin)
{
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::SYNCED) { int rc = mr_job_sync(ctx, nullptr); if (rc) return rc; }
    CU(cudaSetDevice(ctx->cfg.device));
    const uint64_t n = ctx->info.n_keys;
    cudaStream_t s = ctx->stream;
    if (part && n) { CU(cudaMemcpyAsync(part, ctx->rows.part, n * 4, cudaMemcpyDeviceToHost, s)); ctx->stats.d2h_bytes += n * 4; }
    if (count && n) { CU(cudaMemcpyAsync(count, ctx->rows.count, n * 8, cudaMemcpyDeviceToHost, s)); ctx->stats.d2h_bytes += n * 8; }
    if (key_off) { CU(cudaMemcpyAsync(key_off, ctx->rows.key_off, (n + 1) * 8, cudaMemcpyDeviceToHost, s)); ctx->stats.d2h_bytes += (n + 1) * 8; }
    if (key_bytes && ctx->info.key_bytes) {
        CU(cudaMemcpyAsync(key_bytes, ctx->rows.key_bytes, ctx->info.key_bytes, cudaMemcpyDeviceToHost, s));
        ctx->stats.d2h_bytes += ctx->info.key_bytes;
    }
    if (part_begin) {
        CU(cudaMemcpyAsync(part_begin, ctx->rows.part_begin, ((uint64_t)ctx->cfg.num_reducers + 1) * 8, cudaMemcpyDeviceToHost, s));
        ctx->stats.d2h_bytes += ((uint64_t)ctx->cfg.num_reducers + 1) * 8;
    }
    CU(cudaStreamSynchronize(s));
    return MR_OK;
}

int mr_reduce_partition(mr_ctx *ctx, uint32_t part, mr_result **out)
{
    if (!ctx || !out) return MR_ERR_INVALID;
    *out = nullptr;
    if (part >= ctx->cfg.num_reducers) return fail(ctx, MR_ERR_INVALID, "mr_reduce_partition: partition %u out of range", part);
    if (ctx->state != mr_ctx::SYNCED) { int rc = mr_job_sync(ctx, nullptr); if (rc) return rc; }
    CU(cudaSetDevice(ctx->cfg.device));
    if (!ctx->rows_cached) {
        ctx->h_part_begin.assign((size_t)ctx->cfg.num_reducers + 1, 0);
        CU(cudaMemcpyAsync(ctx->h_part_begin.data(), ctx->rows.part_begin, ctx->h_part_begin.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        ctx->stats.d2h_bytes += ctx->h_part_begin.size() * 8;
        ctx->rows_cached = true;
    }
    const uint64_t first = ctx->h_part_begin[part], last = ctx->h_part_begin[part + 1];
    mr_result *r = new mr_result();
    r->ctx = ctx; r->first = first; r->n = last - first; r->cur = 0;
    if (r->n) {
        r->count.resize(r->n); r->key_off.resize(r->n + 1);
        cudaStream_t s = ctx->stream;
        cudaError_t e = cudaMemcpyAsync(r->count.data(), ctx->rows.count + first, r->n * 8, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(r->key_off.data(), ctx->rows.key_off + first, (r->n + 1) * 8, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e == cudaSuccess) {
            uint64_t b0 = r->key_off[0], b1 = r->key_off[r->n];
            r->key_bytes.resize(b1 - b0 + 1);
            if (b1 > b0) e = cudaMemcpyAsync(r->key_bytes.data(), ctx->rows.key_bytes + b0, b1 - b0, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            ctx->stats.d2h_bytes += r->n * 16 + 8 + (b1 - b0);
        }
        if (e != cudaSuccess) { delete r; return fail(ctx, MR_ERR_CUDA, "mr_reduce_partition: %s", cudaGetErrorString(e)); }
    }
    *out = r;
    return MR_OK;
}

int mr_result_next(mr_result *res, const char **key, uint64_t *klen, uint64_t *count)
{
    if (!res) return MR_ERR_INVALID;
    if (res->cur >= res->n) return MR_ERR_END;
    uint64_t i = res->cur++;
    uint64_t b0 = res->key_off[0];
    if (key) *key = (const char *)res->key_bytes.data() + (res->key_off[i] - b0);
    if (klen) *klen = res->key_off[i + 1] - res->key_off[i];
    if (count) *count = res->count[i];
    return MR_OK;
}

void mr_result_release(mr_result *res) { delete res; }

#if defined(MRGPU_SORT_TIMING) || defined(MRGPU_MAP_TIMING) || defined(MRGPU_UNIT_TIMING)
extern "C" const long long *mr_debug_stamps(mr_ctx *ctx) { return ctx->h_ctl->dbg; }
#endif

int mr_stats_get(mr_ctx *ctx, mr_stats *out)
{
    if (!ctx || !out) return MR_ERR_INVALID;
    *out = ctx->stats;
    return MR_OK;
}

// ---------------------------------------------------------------- shuffle
int mr_shuffle_count(mr_ctx *ctx, const uint64_t **dev_counts)
{
    if (!ctx || !dev_counts) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::MAPPED || ctx->imported || ctx->shuffle_counted)
        return fail(ctx, MR_ERR_INVALID, "mr_shuffle_count: call after mr_map_finish, once per job");
    CU(cudaSetDevice(ctx->cfg.device));
    if (ctx->map_dirty) { int rc = mr_map_finish(ctx); if (rc) return rc; }
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    int rc = launch_compact(ctx);
    if (rc) return rc;
    sort_setup_kernel<<<1, 32, 0, s>>>(ctx->d_ctl, ctx->max_keys); // clamps n (its digit plan is redone after import)
    CU(cudaMemsetAsync(ctx->d_shuf, 0, (5ull * W + 1) * 8, s));
    const unsigned g = grid_for(ctx->max_keys, 256, ctx->sm_count * 2);
    export_count_kernel<<<g, 256, 0, s>>>(ctx->ent, W, ctx->d_shuf, ctx->d_shuf + W, ctx->d_ctl);
    ctx->stats.launches += 2;
    CU(cudaGetLastError());
    ctx->shuffle_counted = true;
    *dev_counts = (const uint64_t *)ctx->d_shuf;
    return MR_OK;
}

int mr_shuffle_pack(mr_ctx *ctx, const uint64_t *counts, mr_shuffle_plan *plan)
{
    if (!ctx || !counts || !plan) return MR_ERR_INVALID;
    if (!ctx->shuffle_counted || ctx->imported)
        return fail(ctx, MR_ERR_INVALID, "mr_shuffle_pack: call after mr_shuffle_count");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    unsigned long long *nrec = ctx->d_shuf, *rec_cur = ctx->d_shuf + 2 * W, *byte_cur = ctx->d_shuf + 3 * W,
                       *seg_off = ctx->d_shuf + 4 * W;
    ctx->send_off.assign(W + 1, 0);
    ctx->send_nrec.assign(W, 0);
    uint64_t nkeys = 0;
    for (uint32_t d = 0; d < W; ++d) {
        ctx->send_nrec[d] = counts[d];
        nkeys += counts[d];
        ctx->send_off[d + 1] = ctx->send_off[d] + round_up(counts[d] * REC_BYTES + counts[W + d], 16);
    }
    ctx->info.map_keys = nkeys;
    const uint64_t total = ctx->send_off[W];
    if (total + 64 > ctx->send_cap) {
        CU(cudaStreamSynchronize(s));
        if (ctx->d_send) CU(cudaFree(ctx->d_send));
        ctx->d_send = nullptr;
        ctx->send_cap = total + total / 4 + 4096;
        CU(dmalloc(&ctx->d_send, ctx->send_cap));
    }
    // offsets go up through pinned memory so that no host synchronisation is needed
    unsigned long long *stage = ctx->h_stage;
    for (uint32_t d = 0; d <= W; ++d) stage[d] = ctx->send_off[d];
    CU(cudaMemcpyAsync(seg_off, stage, (W + 1) * 8ull, cudaMemcpyHostToDevice, s));
    const unsigned g = grid_for(ctx->max_keys, 256, ctx->sm_count * 2);
    export_scatter_kernel<<<g, 256, 0, s>>>(ctx->ent, ctx->bytes_base, W, ctx->d_send, seg_off, nrec, rec_cur,
                                            byte_cur, seg_off + 1, ctx->d_ctl);
    ctx->stats.launches++;
    CU(cudaGetLastError());
    plan->send_base = ctx->d_send;
    plan->send_off = ctx->send_off.data();
    plan->send_nrec = ctx->send_nrec.data();
    return MR_OK;
}

int mr_shuffle_pack_fixed(mr_ctx *ctx, uint64_t slot_bytes, const uint8_t **send_base)
{
    if (!ctx || !send_base) return MR_ERR_INVALID;
    if (slot_bytes < 64 || slot_bytes % 16) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_pack_fixed: slot_bytes must be a multiple of 16, >= 64");
    if (!ctx->shuffle_counted) { const uint64_t *dc; int rc = mr_shuffle_count(ctx, &dc); if (rc) return rc; }
    if (ctx->imported) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_pack_fixed: already imported");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    unsigned long long *nrec = ctx->d_shuf, *nbytes = ctx->d_shuf + W, *rec_cur = ctx->d_shuf + 2 * W,
                       *byte_cur = ctx->d_shuf + 3 * W, *seg_off = ctx->d_shuf + 4 * W;
    const uint64_t total = (uint64_t)W * slot_bytes;
    if (total + 64 > ctx->send_cap) {
        CU(cudaStreamSynchronize(s));
        if (ctx->d_send) CU(cudaFree(ctx->d_send));
        ctx->d_send = nullptr;
        ctx->send_cap = total + 4096;
        CU(dmalloc(&ctx->d_send, ctx->send_cap));
    }
    export_header_kernel<<<(W + 255) / 256, 256, 0, s>>>(ctx->d_send, W, slot_bytes, nrec, nbytes, seg_off, ctx->d_ctl);
    const unsigned g = grid_for(ctx->max_keys, 256, ctx->sm_count * 2);
    export_scatter_kernel<<<g, 256, 0, s>>>(ctx->ent, ctx->bytes_base, W, ctx->d_send, seg_off, nrec, rec_cur,
                                            byte_cur, nullptr, ctx->d_ctl);
    ctx->stats.launches += 2;
    CU(cudaGetLastError());
    *send_base = ctx->d_send;
    return MR_OK;
}

int mr_shuffle_import_fixed(mr_ctx *ctx, uint64_t slot_bytes)
{
    if (!ctx) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::MAPPED) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_import_fixed: call after mr_shuffle_pack_fixed");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    if ((uint64_t)W * slot_bytes + 256 > ctx->recv_cap) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_import_fixed: receive buffer smaller than world_size slots");
    CU(cudaMemsetAsync(&ctx->d_ctl->or_lo, 0, offsetof(Ctl, err) - offsetof(Ctl, or_lo), s));
    CU(cudaMemsetAsync(&ctx->d_ctl->n_tie, 0, offsetof(Ctl, tokens) - offsetof(Ctl, n_tie), s));
    ctl_init_kernel<<<1, 1, 0, s>>>(ctx->d_ctl);
    import_fixed_kernel<<<W * IMPORT_CHUNKS, 256, 0, s>>>(ctx->d_recv, W, slot_bytes, ctx->ent, ctx->max_keys, ctx->d_ctl);
    ctx->bytes_base = ctx->d_recv;
    partition_entries_kernel<<<grid_for(ctx->max_keys, 256, ctx->sm_count * 4), 256, 0, s>>>(
        ctx->bytes_base, ctx->cfg.num_reducers, ctx->cfg.partition_kind, ctx->ent, ctx->bufs.item[0], ctx->max_keys, ctx->d_ctl);
    ctx->stats.launches += 3;
    CU(cudaGetLastError());
    ctx->imported = true;
    return MR_OK;
}

int mr_shuffle_recv_buffer(mr_ctx *ctx, uint64_t bytes, void **dev_ptr)
{
    if (!ctx || !dev_ptr) return MR_ERR_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (bytes + ARENA_SLACK >= MAX_BYTE_SPACE) return fail(ctx, MR_ERR_NOMEM, "receive buffer too large");
    if (bytes + 256 > ctx->recv_cap) {
        CU(cudaStreamSynchronize(ctx->stream));
        if (ctx->d_recv) CU(cudaFree(ctx->d_recv));
        ctx->d_recv = nullptr;
        ctx->recv_cap = bytes + bytes / 4 + 4096;
        CU(dmalloc(&ctx->d_recv, ctx->recv_cap));
    }
    CU(cudaMemsetAsync(ctx->d_recv + bytes, 0, 256, ctx->stream));
    *dev_ptr = ctx->d_recv;
    return MR_OK;
}

int mr_shuffle_import(mr_ctx *ctx, const uint64_t *recv_off, const uint64_t *recv_nrec)
{
    if (!ctx || !recv_off || !recv_nrec) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::MAPPED) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_import: call after mr_shuffle_export");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    const uint64_t *nrec = recv_nrec;
    uint64_t base = 0;
    for (uint32_t src = 0; src < W; ++src) {
        uint64_t n = nrec[src];
        if (recv_off[src] + n * REC_BYTES > recv_off[src + 1]) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_import: segment %u too small", src);
        if (base + n > ctx->max_keys) return fail(ctx, MR_ERR_NOMEM, "mr_shuffle_import: more than max_keys records");
        if (n) {
            import_segment_kernel<<<grid_for(n, 256, ctx->sm_count * 4), 256, 0, s>>>(
                ctx->d_recv, recv_off[src], n, recv_off[src + 1] - recv_off[src] - n * REC_BYTES, base, ctx->ent, ctx->d_ctl);
            ctx->stats.launches++;
        }
        base += n;
    }
    // fresh control block for the reduce side, keeping the token count and errors
    CU(cudaMemsetAsync(&ctx->d_ctl->or_lo, 0, offsetof(Ctl, err) - offsetof(Ctl, or_lo), s));
    CU(cudaMemsetAsync(&ctx->d_ctl->n_tie, 0, offsetof(Ctl, tokens) - offsetof(Ctl, n_tie), s));
    ctl_init_kernel<<<1, 1, 0, s>>>(ctx->d_ctl);
    ctl_set_n_kernel<<<1, 1, 0, s>>>(ctx->d_ctl, base);
    ctx->bytes_base = ctx->d_recv;
    partition_entries_kernel<<<grid_for(std::max<uint64_t>(base, 1), 256, ctx->sm_count * 4), 256, 0, s>>>(
        ctx->bytes_base, ctx->cfg.num_reducers, ctx->cfg.partition_kind, ctx->ent, ctx->bufs.item[0], ctx->max_keys, ctx->d_ctl);
    ctx->stats.launches += 3;
    CU(cudaGetLastError());
    ctx->imported = true;
    return MR_OK;
}


// ------------------------------------------------- the library's own transport (NCCL)
// A host without torch.distributed (the reference's Lua worker) still needs the all-to-all
// between mr_shuffle_pack_fixed and mr_shuffle_import_fixed.  libnccl.so.2 is opened at run
// time (no link-time dependency; the copy a host process already loaded is reused) and only
// the point-to-point and all-reduce calls are bound.
namespace {

struct NcclId { char internal[MR_NCCL_ID_BYTES]; };
typedef void *NcclComm;
enum { NCCL_UINT8 = 1, NCCL_UINT64 = 5, NCCL_MAX = 2 }; // ncclDataType_t / ncclRedOp_t values of NCCL 2.x

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string why;
};

NcclApi g_nccl;
std::once_flag g_nccl_once;

void nccl_open()
{
    const char *names[] = { getenv("MRGPU_NCCL_LIB"), "libnccl.so.2", "libnccl.so" };
    for (const char *n : names) {
        if (!n || !*n) continue;
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
        g_nccl.why = dlerror();
    }
    if (!g_nccl.lib) return;
    struct { const char *name; void **slot; } syms[] = {
        { "ncclGetUniqueId", (void **)&g_nccl.GetUniqueId },   { "ncclCommInitRank", (void **)&g_nccl.CommInitRank },
        { "ncclCommDestroy", (void **)&g_nccl.CommDestroy },   { "ncclGroupStart", (void **)&g_nccl.GroupStart },
        { "ncclGroupEnd", (void **)&g_nccl.GroupEnd },         { "ncclSend", (void **)&g_nccl.Send },
        { "ncclRecv", (void **)&g_nccl.Recv },                 { "ncclAllReduce", (void **)&g_nccl.AllReduce },
        { "ncclGetErrorString", (void **)&g_nccl.GetErrorString },
    };
    for (auto &sy : syms) {
        *sy.slot = dlsym(g_nccl.lib, sy.name);
        if (!*sy.slot) {
            g_nccl.why = std::string("libnccl lacks ") + sy.name;
            g_nccl.lib = nullptr;
            return;
        }
    }
}

const NcclApi *nccl_api()
{
    std::call_once(g_nccl_once, nccl_open);
    return g_nccl.lib ? &g_nccl : nullptr;
}

#define NC(call)                                                                                   \
    do {                                                                                           \
        int r_ = (call);                                                                           \
        if (r_ != 0)                                                                               \
            return fail(ctx, MR_ERR_CUDA, "%s failed: %s", #call, g_nccl.GetErrorString(r_));     \
    } while (0)

void nccl_release(mr_ctx *ctx)
{
    if (ctx->nccl_comm && g_nccl.lib) g_nccl.CommDestroy((NcclComm)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
}

// largest outgoing slot of this rank: header + records + long-key bytes, rounded as the packer does
__global__ void segment_max_kernel(const unsigned long long *nrec, const unsigned long long *nbytes, uint32_t W,
                                   unsigned long long *out)
{
    unsigned long long m = 0;
    for (uint32_t d = threadIdx.x; d < W; d += 32) {
        unsigned long long b = 16 + nrec[d] * REC_BYTES + nbytes[d];
        m = max(m, (b + 15) / 16 * 16);
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(~0u, m, o));
    if (threadIdx.x == 0) *out = m;
}

} // namespace

int mr_nccl_unique_id(uint8_t *id)
{
    mr_ctx *ctx = nullptr;
    if (!id) return MR_ERR_INVALID;
    const NcclApi *api = nccl_api();
    if (!api) return fail(ctx, MR_ERR_CUDA, "libnccl.so.2 not loadable: %s", g_nccl.why.c_str());
    NcclId nid;
    NC(api->GetUniqueId(&nid));
    memcpy(id, nid.internal, MR_NCCL_ID_BYTES);
    return MR_OK;
}

int mr_nccl_init(mr_ctx *ctx, const uint8_t *id)
{
    if (!ctx || !id) return MR_ERR_INVALID;
    const NcclApi *api = nccl_api();
    if (!api) return fail(ctx, MR_ERR_CUDA, "libnccl.so.2 not loadable: %s", g_nccl.why.c_str());
    if (ctx->nccl_comm) return fail(ctx, MR_ERR_INVALID, "mr_nccl_init: context already has a communicator");
    CU(cudaSetDevice(ctx->cfg.device));
    NcclId nid;
    memcpy(nid.internal, id, MR_NCCL_ID_BYTES);
    NcclComm comm = nullptr;
    NC(api->CommInitRank(&comm, (int)ctx->cfg.world_size, nid, (int)ctx->cfg.rank));
    ctx->nccl_comm = comm;
    return MR_OK;
}

int mr_shuffle_nccl(mr_ctx *ctx, uint64_t slot_bytes, uint64_t *slot_used)
{
    if (!ctx) return MR_ERR_INVALID;
    if (!ctx->nccl_comm) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_nccl: call mr_nccl_init first");
    const NcclApi *api = nccl_api();
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    NcclComm comm = (NcclComm)ctx->nccl_comm;
    if (slot_bytes == 0) {
        // size the slots from this job: largest segment over all ranks, plus a quarter (one host sync)
        const uint64_t *dc;
        if (!ctx->shuffle_counted) { int rc = mr_shuffle_count(ctx, &dc); if (rc) return rc; }
        unsigned long long *d_max = ctx->d_shuf + 5ull * W + 1;
        segment_max_kernel<<<1, 32, 0, s>>>(ctx->d_shuf, ctx->d_shuf + W, W, d_max);
        ctx->stats.launches++;
        NC(api->AllReduce(d_max, d_max, 1, NCCL_UINT64, NCCL_MAX, comm, s));
        CU(cudaMemcpyAsync(ctx->h_stage, d_max, 8, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        const uint64_t m = ctx->h_stage[0];
        slot_bytes = round_up(m + m / 4 + 16, 4096);
    }
    const uint8_t *send = nullptr;
    void *recv = nullptr;
    int rc = mr_shuffle_pack_fixed(ctx, slot_bytes, &send);
    if (rc) return rc;
    rc = mr_shuffle_recv_buffer(ctx, (uint64_t)W * slot_bytes, &recv);
    if (rc) return rc;
    NC(api->GroupStart());
    for (uint32_t p = 0; p < W; ++p) {
        NC(api->Send(send + (uint64_t)p * slot_bytes, slot_bytes, NCCL_UINT8, (int)p, comm, s));
        NC(api->Recv((uint8_t *)recv + (uint64_t)p * slot_bytes, slot_bytes, NCCL_UINT8, (int)p, comm, s));
    }
    NC(api->GroupEnd());
    if (slot_used) *slot_used = slot_bytes;
    return mr_shuffle_import_fixed(ctx, slot_bytes);
}

// ------------------------------------------------- peer-to-peer shuffle (p2p_kernels.cuh)
namespace {

void p2p_release(mr_ctx *ctx)
{
    for (size_t r = 0; r < ctx->p2p_peer.size(); ++r)
        if (ctx->p2p_ipc[r] && ctx->p2p_peer[r]) cudaIpcCloseMemHandle(ctx->p2p_peer[r]);
    ctx->p2p_peer.clear();
    ctx->p2p_ipc.clear();
    cudaFree(ctx->p2p_win); ctx->p2p_win = nullptr;
    cudaFree(ctx->d_p2p_buf); ctx->d_p2p_buf = nullptr;
    cudaFree(ctx->d_p2p_win); ctx->d_p2p_win = nullptr;
    cudaFree(ctx->d_p2p_done); ctx->d_p2p_done = nullptr;
    ctx->p2p_attached = false;
    ctx->p2p_slot = 0;
}

} // namespace

static_assert(sizeof(mr_peer_info) == 128, "mr_peer_info is part of the ABI");
static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(((mr_peer_info *)0)->ipc_handle), "IPC handle size");

int mr_p2p_setup(mr_ctx *ctx, uint64_t slot_bytes, mr_peer_info *mine)
{
    if (!ctx || !mine) return MR_ERR_INVALID;
    if (slot_bytes < 64 || slot_bytes % 32) return fail(ctx, MR_ERR_INVALID, "mr_p2p_setup: slot_bytes must be a multiple of 32, >= 64");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    if (2ull * W * slot_bytes + ARENA_SLACK >= MAX_BYTE_SPACE) return fail(ctx, MR_ERR_NOMEM, "mr_p2p_setup: window too large");
    CU(cudaStreamSynchronize(ctx->stream));
    p2p_release(ctx);
    ctx->p2p_slot = slot_bytes;
    ctx->p2p_win_bytes = P2P_FLAG_BYTES + 2ull * W * slot_bytes + 256;
    CU(cudaMalloc((void **)&ctx->p2p_win, ctx->p2p_win_bytes));
    CU(cudaMemset(ctx->p2p_win, 0, P2P_FLAG_BYTES)); // flags: no epoch seen yet
    CU(dmalloc(&ctx->d_p2p_buf, 2ull * W));
    CU(dmalloc(&ctx->d_p2p_win, W));
    CU(dmalloc(&ctx->d_p2p_done, 1));
    CU(cudaMemset(ctx->d_p2p_done, 0, sizeof(unsigned int)));
    CU(cudaStreamSynchronize(0)); // both clears have happened before the handle leaves this rank
    ctx->p2p_epoch = 0;
    memset(mine, 0, sizeof *mine);
    mine->pid = (uint64_t)getpid();
    mine->dev_ptr = (uint64_t)(uintptr_t)ctx->p2p_win;
    mine->win_bytes = ctx->p2p_win_bytes;
    mine->slot_bytes = slot_bytes;
    mine->device = ctx->cfg.device;
    mine->rank = ctx->cfg.rank;
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, ctx->p2p_win));
    memcpy(mine->ipc_handle, &h, sizeof h);
    return MR_OK;
}

int mr_p2p_attach(mr_ctx *ctx, const mr_peer_info *all)
{
    if (!ctx || !all) return MR_ERR_INVALID;
    if (!ctx->p2p_win) return fail(ctx, MR_ERR_INVALID, "mr_p2p_attach: call mr_p2p_setup first");
    if (ctx->p2p_attached) return fail(ctx, MR_ERR_INVALID, "mr_p2p_attach: already attached (mr_p2p_setup starts over)");
    CU(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->cfg.world_size;
    ctx->p2p_peer.assign(W, nullptr);
    ctx->p2p_ipc.assign(W, 0);
    for (uint32_t r = 0; r < W; ++r) {
        const mr_peer_info &pi = all[r];
        if (pi.rank != r) return fail(ctx, MR_ERR_INVALID, "mr_p2p_attach: entry %u describes rank %u", r, pi.rank);
        if (pi.slot_bytes != ctx->p2p_slot)
            return fail(ctx, MR_ERR_INVALID, "mr_p2p_attach: rank %u uses slot_bytes %llu, this rank %llu", r,
                        (unsigned long long)pi.slot_bytes, (unsigned long long)ctx->p2p_slot);
        if (r == ctx->cfg.rank) {
            if (pi.dev_ptr != (uint64_t)(uintptr_t)ctx->p2p_win) return fail(ctx, MR_ERR_INVALID, "mr_p2p_attach: entry %u is not this context's window", r);
            ctx->p2p_peer[r] = ctx->p2p_win;
        } else if (pi.pid == (uint64_t)getpid()) {
            // another context of this process: the pointer is valid here, the device must be reachable
            if (pi.device != ctx->cfg.device) {
                int can = 0;
                CU(cudaDeviceCanAccessPeer(&can, ctx->cfg.device, pi.device));
                if (!can) return fail(ctx, MR_ERR_UNSUPPORTED, "mr_p2p_attach: device %d cannot reach device %d", ctx->cfg.device, pi.device);
                cudaError_t e = cudaDeviceEnablePeerAccess(pi.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) return fail(ctx, MR_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", pi.device, cudaGetErrorString(e));
            }
            ctx->p2p_peer[r] = (uint8_t *)(uintptr_t)pi.dev_ptr;
        } else {
            cudaIpcMemHandle_t h;
            memcpy(&h, pi.ipc_handle, sizeof h);
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return fail(ctx, MR_ERR_UNSUPPORTED, "mr_p2p_attach: cannot map the window of rank %u (%s); use mr_shuffle_nccl", r, cudaGetErrorString(e));
            }
            ctx->p2p_peer[r] = (uint8_t *)ptr;
            ctx->p2p_ipc[r] = 1;
        }
    }
    std::vector<uint8_t *> tab(3ull * W);
    for (uint32_t r = 0; r < W; ++r) {
        tab[r] = ctx->p2p_peer[r] + P2P_FLAG_BYTES;                                          // parity 0
        tab[W + r] = ctx->p2p_peer[r] + P2P_FLAG_BYTES + (uint64_t)W * ctx->p2p_slot;        // parity 1
        tab[2ull * W + r] = ctx->p2p_peer[r];
    }
    CU(cudaMemcpy(ctx->d_p2p_buf, tab.data(), 2ull * W * sizeof(uint8_t *), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(ctx->d_p2p_win, tab.data() + 2ull * W, W * sizeof(uint8_t *), cudaMemcpyHostToDevice));
    ctx->p2p_attached = true;
    return MR_OK;
}

int mr_shuffle_p2p(mr_ctx *ctx)
{
    if (!ctx) return MR_ERR_INVALID;
    if (!ctx->p2p_attached) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_p2p: call mr_p2p_setup and mr_p2p_attach first");
    if (ctx->state != mr_ctx::MAPPED || ctx->imported || ctx->shuffle_counted)
        return fail(ctx, MR_ERR_INVALID, "mr_shuffle_p2p: call after mr_map_finish, once per job");
    CU(cudaSetDevice(ctx->cfg.device));
    if (ctx->map_dirty) { int rc = mr_map_finish(ctx); if (rc) return rc; }
    const uint32_t W = ctx->cfg.world_size;
    cudaStream_t s = ctx->stream;
    // sender: ONE kernel from the global table to records in the owners' slots, headers and flags
    CU(cudaMemsetAsync(ctx->d_shuf, 0, (5ull * W + 1) * 8, s));
    unsigned long long *rec_cur = ctx->d_shuf + 2 * W, *byte_cur = ctx->d_shuf + 3 * W;
    const uint64_t epoch = ++ctx->p2p_epoch;
    const uint32_t parity = (uint32_t)(epoch & 1);
    P2PParams p;
    p.peer_buf = ctx->d_p2p_buf + (uint64_t)parity * W;
    p.peer_win = ctx->d_p2p_win;
    p.world = W;
    p.rank = ctx->cfg.rank;
    p.slot_bytes = ctx->p2p_slot;
    p.epoch = epoch;
    const uint64_t tiles = (ctx->slots + 256 * P2P_SEND_ITEMS - 1) / (256 * P2P_SEND_ITEMS);
    unsigned per_sm;
    table_reader_config(ctx, 1, (const void *)p2p_send_kernel, &per_sm);
    const unsigned g = (unsigned)std::min<uint64_t>(tiles, (uint64_t)ctx->sm_count * per_sm);
    const size_t dyn = 24ull * W;
    if (dyn > 8192) CU(cudaFuncSetAttribute(p2p_send_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    p2p_send_kernel<<<g, 256, dyn, s>>>(p, ctx->d_gtab, ctx->slots, ctx->bytes_base, ctx->cfg.num_reducers, ctx->cfg.partition_kind,
                                      ctx->owners(), rec_cur, byte_cur, ctx->d_p2p_done, ctx->d_ctl);
    ctx->table_dirty = false;
    // owner: a fresh control block (token count and errors kept), then index the window in place
    uint8_t *mybuf = ctx->p2p_win + P2P_FLAG_BYTES + (uint64_t)parity * W * ctx->p2p_slot;
    ctl_owner_reset_kernel<<<1, 1, 0, s>>>(ctx->d_ctl);
    const uint32_t chunks = std::max<uint32_t>(1, std::min<uint32_t>(P2P_MAX_CHUNKS, (uint32_t)ctx->sm_count * 4 / W));
    p2p_index_kernel<<<W * chunks, 256, 0, s>>>(mybuf, (const unsigned long long *)ctx->p2p_win, W, chunks, ctx->p2p_slot, epoch,
                                                ctx->bufs.item[0], ctx->max_keys, ctx->d_ctl);
    ctx->stats.launches += 3;
    CU(cudaGetLastError());
    ctx->p2p_sent_pending = true;
    ctx->ent.e = (GEntry *)mybuf;
    ctx->bytes_base = mybuf;
    ctx->imported = true;
    ctx->shuffle_counted = true;
    return MR_OK;
}

// ------------------------------------------------- owner ranks by key range
namespace {

// sort word of round 0 on the host: same value as sort_chunk(key, bytes, 0) in common.cuh
uint64_t host_sort_word(const uint8_t *key, uint64_t len)
{
    uint64_t w = 0;
    for (uint64_t i = 0; i < 7; ++i) w = (w << 8) | (i < len ? key[i] : 0);
    return (w << 8) | (len < 8 ? len : 8);
}

} // namespace

int mr_shuffle_set_owners(mr_ctx *ctx, uint32_t n, const uint32_t *part, const uint64_t *word)
{
    if (!ctx) return MR_ERR_INVALID;
    const uint32_t W = ctx->cfg.world_size;
    if (n != 0 && n != W - 1) return fail(ctx, MR_ERR_INVALID, "mr_shuffle_set_owners: need 0 or world_size-1 = %u splitters, got %u", W - 1, n);
    if (n && (!part || !word)) return MR_ERR_INVALID;
    for (uint32_t i = 1; i < n; ++i)
        if (part[i] < part[i - 1] || (part[i] == part[i - 1] && word[i] < word[i - 1]))
            return fail(ctx, MR_ERR_INVALID, "mr_shuffle_set_owners: splitters must ascend");
    if (ctx->state == mr_ctx::MAPPED && (ctx->shuffle_counted || ctx->imported))
        return fail(ctx, MR_ERR_INVALID, "mr_shuffle_set_owners: a shuffle is under way");
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaStreamSynchronize(ctx->stream)); // earlier jobs may still read the old splitters
    ctx->split_part.assign(part, part + n);
    ctx->split_word.assign(word, word + n);
    ctx->lsd_jobs_left = 0;   // other keys will arrive here from now on: the shared-memory reduce path gets a fresh try
    if (n) {
        CU(cudaMemcpy(ctx->d_split_part, part, n * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(ctx->d_split_word, word, n * sizeof(uint64_t), cudaMemcpyHostToDevice));
    }
    return MR_OK;
}

int mr_owner_of(const mr_ctx *ctx, const void *key, uint64_t klen, uint32_t *owner)
{
    if (!ctx || !owner || (klen && !key)) return MR_ERR_INVALID;
    uint32_t part;
    if (ctx->cfg.partition_kind == MR_PARTITION_MOD_U32) { // number keys travel as 4 big-endian bytes
        if (klen != 4) return MR_ERR_INVALID;
        const uint8_t *b = (const uint8_t *)key;
        part = ((uint32_t)b[0] << 24 | (uint32_t)b[1] << 16 | (uint32_t)b[2] << 8 | b[3]) % ctx->cfg.num_reducers;
    } else {
        part = mr_partition_of(key, klen, ctx->cfg.num_reducers);
    }
    if (ctx->split_word.empty()) { *owner = part % ctx->cfg.world_size; return MR_OK; }
    const uint64_t w = host_sort_word((const uint8_t *)key, klen);
    uint32_t r = 0;
    while (r < ctx->split_word.size() &&
           (ctx->split_part[r] < part || (ctx->split_part[r] == part && ctx->split_word[r] <= w))) ++r;
    *owner = r;
    return MR_OK;
}

int mr_row_sort_word(mr_ctx *ctx, uint64_t row, uint32_t *part, uint64_t *word)
{
    if (!ctx || !part || !word) return MR_ERR_INVALID;
    if (ctx->state != mr_ctx::SYNCED) { int rc = mr_job_sync(ctx, nullptr); if (rc) return rc; }
    if (row >= ctx->info.n_keys) return fail(ctx, MR_ERR_INVALID, "mr_row_sort_word: row %llu of %llu", (unsigned long long)row,
                                             (unsigned long long)ctx->info.n_keys);
    CU(cudaSetDevice(ctx->cfg.device));
    uint64_t off[2];
    uint8_t kb[8] = { 0 };
    CU(cudaMemcpyAsync(off, ctx->rows.key_off + row, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(part, ctx->rows.part + row, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint64_t len = off[1] - off[0];
    if (len) CU(cudaMemcpyAsync(kb, ctx->rows.key_bytes + off[0], len < 7 ? len : 7, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    *word = host_sort_word(kb, len);
    return MR_OK;
}
