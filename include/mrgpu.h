/*
 * mrgpu.h -- C ABI of the B200 data plane that stands behind lua-mapreduce's
 * task-module API (map -> combine -> partition -> shuffle -> sort/merge -> reduce).
 *
 * The reference has no FFI: its data plane is Lua (mapreduce/job.lua) called from
 * mapreduce/worker.lua:73.  Each entry point below names the reference code it
 * replaces (paths relative to the reference root); INTEGRATION.md shows the Lua
 * C-API / LuaJIT-FFI binding a maintainer would add at those seams.
 *
 * Conventions: plain C types only; every function returns MR_OK (0) or an error
 * code, with the text available from mr_last_error(); no exceptions or longjmp
 * cross the boundary, so Lua glue turns a non-zero return into error(msg) and the
 * reference's BROKEN/FAILED retry machinery (job.lua:322-342, worker.lua:116-137)
 * is unchanged.  One context per process per GPU; calls on a context must be
 * serialised by the caller (a Lua worker is single-threaded, worker.lua:54-86).
 * Input buffers are borrowed only for the duration of the call.
 */
#ifndef MRGPU_H
#define MRGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MRGPU_ABI_VERSION 1

typedef struct mr_ctx mr_ctx;
typedef struct mr_result mr_result;

enum mr_status {
    MR_OK = 0,
    MR_ERR_INVALID = 1,     /* bad argument or call order */
    MR_ERR_CUDA = 2,        /* CUDA runtime error (sticky: destroy and recreate the context) */
    MR_ERR_NOMEM = 3,       /* arena / key capacity exceeded */
    MR_ERR_OVERFLOW = 4,    /* hash table full or a token longer than MR_MAX_TOKEN_BYTES */
    MR_ERR_UNSUPPORTED = 5, /* task module asks for a map/partition/reduce kind not built */
    MR_ERR_END = 6          /* mr_result_next: no more rows */
};

/* What the task module's `native` descriptor may name.  The Lua functions stay the
 * semantic definition; a module without a descriptor is simply not accelerated. */
enum mr_map_kind {
    MR_MAP_TOKENIZE_WS = 0, /* examples/WordCount/mapfn.lua:3-9: emit(word, 1) per whitespace-separated token */
    MR_MAP_RECORDS_U32 = 1, /* a mapfn that emits (number key, number value) records: emit, job.lua:83-97 */
    MR_MAP_BIGRAM_WS = 2    /* mapfn.lua:3-9 emitting (w_i .. " " .. w_{i+1}, 1) for neighbours inside a line */
};
enum mr_reduce_kind {       /* reducefn = combinerfn; all associative, commutative, idempotent (job.lua:104-106).  The
                             * reference calls such a reducer only for keys with more than one value and otherwise
                             * always (job.lua:264-284, examples/WordCount/reducefn2.lua); for these four the rows
                             * are the same either way. */
    MR_REDUCE_SUM_U64 = 0,  /* examples/WordCount/reducefn.lua:1-5: emit(sum of values) */
    MR_REDUCE_MIN_U64 = 1,  /* emit(math.min(unpack(values))): number keys (mr_map_records) or string keys with values (mr_map_pairs) */
    MR_REDUCE_MAX_U64 = 2,  /* emit(math.max(unpack(values))): the same */
    MR_REDUCE_COUNT = 3     /* emit(number of values): records, or tokenised text (where it equals the sum of the emitted ones) */
};
enum mr_partition_kind {
    MR_PARTITION_FNV1_LUA52 = 0, /* examples/WordCount/partitionfn.lua:8-16 (string keys) */
    MR_PARTITION_MOD_U32 = 1     /* partitionfn(key) = key % NUM_REDUCERS (number keys; job.lua:203-207) */
};
/* Supported pairs: (TOKENIZE_WS | BIGRAM_WS, FNV1_LUA52) and (RECORDS_U32, MOD_U32).  Number keys are ordered
 * numerically (Lua `<`, utils.lua:214) and come back in the rows as 4 big-endian bytes. */

#define MR_MAX_TOKEN_BYTES ((1u << 26) - 1)

typedef struct mr_config {
    uint32_t struct_size;   /* sizeof(mr_config), for ABI growth */
    int32_t device;         /* CUDA device ordinal */
    uint32_t num_reducers;  /* NUM_REDUCERS of partitionfn.lua:2 (15 in-tree, 10 in README.md:59) */
    uint32_t map_kind, reduce_kind, partition_kind;
    uint64_t arena_bytes;   /* capacity for the input splits of one job, resident in HBM */
    uint64_t max_keys;      /* capacity for distinct keys of one job (0 = derive from arena) */
    uint64_t table_slots;   /* global hash-table slots, power of two (0 = 2 x max_keys rounded up) */
    void *stream;           /* cudaStream_t to run on, NULL = a private stream */
    uint32_t world_size;    /* ranks taking part in the shuffle (1 = no shuffle) */
    uint32_t rank;
} mr_config;

typedef struct mr_job_info {
    uint64_t n_keys;        /* rows this rank holds after reduce (distinct keys it owns) */
    uint64_t n_tokens;      /* emits seen by this rank's map side */
    uint64_t input_bytes;   /* split bytes mapped by this rank (padding excluded) */
    uint64_t key_bytes;     /* sum of len(key) over this rank's rows */
    uint64_t map_keys;      /* distinct keys after this rank's combine, before the shuffle */
    uint32_t tie_rounds;    /* extra ordering rounds needed for keys equal in their first 7 bytes */
    uint32_t flags;         /* MR_JOB_* bits: which sort path reduced the job (both give the same rows) */
} mr_job_info;
#define MR_JOB_REDUCED_LSD 1u   /* rows came from the global LSD radix path (reduce_kernels.cuh) */
#define MR_JOB_MSD_DECLINED 2u  /* the shared-memory path (msd_kernels.cuh) was tried first and declined; bits 8.. say why */

/* Device-side layout of the reduced rows, in the reference's final order
 * (partition ascending, key ascending: server.lua:365, utils.lua:214).  All
 * pointers are device pointers owned by the context, valid until mr_job_begin. */
typedef struct mr_device_rows {
    uint64_t n;
    const uint32_t *part;      /* [n] */
    const uint64_t *count;     /* [n] */
    const uint64_t *key_off;   /* [n+1] */
    const uint8_t *key_bytes;  /* [key_off[n]] */
    const uint64_t *part_begin; /* [num_reducers+1] row index where each partition starts */
} mr_device_rows;

typedef struct mr_stats {
    float map_ms;           /* CUDA-event time of the last mr_map_finish */
    float reduce_ms;        /* CUDA-event time of the last mr_reduce_all */
    uint64_t h2d_bytes;     /* bytes copied host->device by mr_map_split since mr_job_begin */
    uint64_t d2h_bytes;     /* bytes copied device->host by result calls since mr_job_begin */
    uint64_t launches;      /* kernels of this library launched since mr_job_begin */
} mr_stats;

/* Context = what one worker process keeps between jobs (job.lua:64-65 caches its
 * modules per process the same way).  Allocates all device memory up front. */
int mr_ctx_create(const mr_config *cfg, mr_ctx **out);
void mr_ctx_destroy(mr_ctx *ctx);
/* Text of the last error on ctx; ctx == NULL gives the last mr_ctx_create failure. */
const char *mr_last_error(const mr_ctx *ctx);
int mr_abi_version(void);

/* Start a job (task iteration).  keep_input != 0 keeps the splits already staged
 * in HBM (the "loop" iteration of server.lua:386-402 with map affinity,
 * task.lua:279-293) and only clears keys and results. */
int mr_job_begin(mr_ctx *ctx, int keep_input);

/* Map job for one split: replaces the closure of job.lua:166-227 for
 * MR_MAP_TOKENIZE_WS, i.e. mapfn.lua:3-9 + emit (job.lua:83-97).  The bytes are
 * the content of the file the reference's taskfn names (taskfn.lua:8-11); they
 * are staged into the HBM arena by a host->device copy the call queues.  A
 * pageable buffer may be reused as soon as the call returns; a page-locked
 * (pinned) buffer is read by the copy engine asynchronously and must stay
 * unchanged until mr_map_finish has returned and the stream has caught up
 * (mr_job_sync, or any later result call).  Tokens never span splits. */
int mr_map_split(mr_ctx *ctx, uint64_t split_id, const void *bytes, uint64_t n);
/* Same, the library reading the file itself (mapfn.lua:4 opens its value with io.lines):
 * one page-locked buffer per context, so consecutive files do not overlap their copies;
 * a host that wants the overlap reads the files itself and calls mr_map_split. */
int mr_map_split_path(mr_ctx *ctx, uint64_t split_id, const char *path);
/* Same with bytes already in device memory (device->device copy). */
int mr_map_split_device(mr_ctx *ctx, uint64_t split_id, const void *dev_bytes, uint64_t n);

/* Map job whose output is already (key, value) number records (MR_MAP_RECORDS_U32): what emit
 * (job.lua:83-97) would receive from a mapfn that emits numbers, n records at a time.  The values of
 * equal keys are summed (reducefn.lua:1-5).  Host pointers: host->device copy inside the call. */
int mr_map_records(mr_ctx *ctx, uint64_t split_id, const uint32_t *keys, const uint32_t *values, uint64_t n);
int mr_map_records_device(mr_ctx *ctx, uint64_t split_id, const uint32_t *dev_keys,
                          const uint32_t *dev_values, uint64_t n);

/* Reduce-side input from run files: what the reduce job reads from the runs other mappers wrote
 * (job.lua:253-265; each line `return <key>,{<values>}` is one load(line)(), utils.lua:218-230, and
 * merge_min_keys concatenates the value lists of equal keys, utils.lua:232-247).  n string keys given
 * as bytes + offsets (key_off[0] = 0, key i = key_bytes[key_off[i] .. key_off[i+1]), 1 byte or more)
 * and one value per key -- the host reduces a line's value list with the job's reducer before the
 * call (a sum for MR_REDUCE_SUM_U64).  The pairs are combined into the same table as the text
 * splits of this job, so GPU-mapped splits and runs written by reference mappers merge in one
 * mr_reduce_all.  MR_MAP_TOKENIZE_WS contexts, reducers SUM / MIN / MAX; host pointers, borrowed
 * for the call (it returns after the copy).  n_tokens of the job counts one per pair.  A context
 * that ingested pairs cannot mr_job_begin(keep_input). */
int mr_map_pairs(mr_ctx *ctx, uint64_t split_id, const void *key_bytes, const uint64_t *key_off,
                 const uint64_t *values, uint64_t n);

/* Runs tokenise + hash-combine over every split staged since the last
 * mr_map_finish, then the partitioner on the distinct keys: job.lua:182-215
 * (emit, combiner, partitionfn) for all staged map jobs at once.  Asynchronous. */
int mr_map_finish(mr_ctx *ctx);

/* Sort/merge + reduce of every partition this rank owns: replaces
 * utils.keys_sorted (utils.lua:123-128), utils.merge_iterator + heap
 * (utils.lua:206-271, heap.lua) and the reduce loop job.lua:264-284.
 * Asynchronous; rows stay in HBM. */
int mr_reduce_all(mr_ctx *ctx);

/* Waits for the job's device work, finishes rare slow paths, fills info. */
int mr_job_sync(mr_ctx *ctx, mr_job_info *info);

/* Device view of the rows (after mr_job_sync). */
int mr_device_rows_get(mr_ctx *ctx, mr_device_rows *rows);

/* Bulk device->host copy of all rows (after mr_job_sync).  Any pointer may be
 * NULL to skip that array.  key_bytes must hold info.key_bytes bytes. */
int mr_rows_fetch(mr_ctx *ctx, uint32_t *part, uint64_t *count, uint64_t *key_off,
                  uint8_t *key_bytes, uint64_t *part_begin);

/* Result cursor of one partition = reading the file result.P<part> the reduce job
 * writes (job.lua:272-273,287) the way server_final does (server.lua:358-383):
 * keys ascending.  A partition that received no key yields an empty cursor (the
 * reference creates no file for it, server.lua:297-324).  key stays valid until
 * mr_result_release. */
int mr_reduce_partition(mr_ctx *ctx, uint32_t part, mr_result **out);
int mr_result_next(mr_result *res, const char **key, uint64_t *klen, uint64_t *count);
void mr_result_release(mr_result *res);

/* Shuffle seam (job.lua:208-221 write, job.lua:253-260 read).  After mr_map_finish:
 *   1. mr_shuffle_count   tallies this rank's combined (key, count) records per owner rank
 *                         (owner = partition mod world_size); *dev_counts is a DEVICE array
 *                         [2 * world_size]: records per destination, then long-key bytes per
 *                         destination.  Asynchronous: the caller gathers the counts of all ranks
 *                         (one all-gather) and reads them once.
 *   2. mr_shuffle_pack    given the HOST copy of this rank's counts, packs one segment per
 *                         destination: 24-byte records {k0, k1, count} followed by the bytes of
 *                         keys longer than 15 bytes; segment size = (24 * records + bytes)
 *                         rounded up to 16.  Asynchronous.
 *   3. the caller moves the segments (NCCL all-to-all) into mr_shuffle_recv_buffer and calls
 *      mr_shuffle_import before mr_reduce_all.
 * Pointers in the plan are owned by the context.
 * Stream contract: every call of this seam only QUEUES work on the context's stream (cfg.stream).
 * The caller's reads of *dev_counts, its transfer out of send_base and into the receive buffer must
 * be ordered after the preceding call and before the following one ON THAT STREAM: run the
 * collective on the same stream (create the context on the collective library's stream), or put
 * events between the two streams.  A context created with stream = NULL owns a private stream the
 * caller cannot name, so it cannot be used with a caller-run collective (mr_shuffle_nccl and
 * mr_shuffle_p2p do their own ordering).  mr_shuffle_pack checks on the device that the counts it
 * was given cover what the count kernel tallied; counts that are too small surface as
 * MR_ERR_OVERFLOW from mr_job_sync and nothing is written past a segment. */
typedef struct mr_shuffle_plan {
    const uint8_t *send_base;  /* device buffer holding every outgoing segment */
    const uint64_t *send_off;  /* host array [world_size+1]: byte offsets per destination */
    const uint64_t *send_nrec; /* host array [world_size]: records per destination */
} mr_shuffle_plan;
int mr_shuffle_count(mr_ctx *ctx, const uint64_t **dev_counts);
int mr_shuffle_pack(mr_ctx *ctx, const uint64_t *counts, mr_shuffle_plan *plan);
/* Fixed-slot variant with no host synchronisation: every destination's segment goes into a slot
 * of slot_bytes (multiple of 16) led by a 16-byte header {records, long-key bytes}; the caller runs
 * a plain equal-split all-to-all of world_size x slot_bytes into mr_shuffle_recv_buffer and calls
 * mr_shuffle_import_fixed.  A segment that does not fit is reported by mr_job_sync as
 * MR_ERR_OVERFLOW (rerun the job with a larger slot or the exact-size path above). */
int mr_shuffle_pack_fixed(mr_ctx *ctx, uint64_t slot_bytes, const uint8_t **send_base);
int mr_shuffle_import_fixed(mr_ctx *ctx, uint64_t slot_bytes);

/* The library's own transport for that all-to-all, for a host that has no collective library of
 * its own (the reference's Lua worker: its shuffle medium is GridFS, job.lua:208-221 and
 * utils.lua:176-186; here it is NCCL over NVLink).  libnccl.so.2 is opened at run time
 * (override the path with MRGPU_NCCL_LIB); nothing else in the library needs it.
 *   mr_nccl_unique_id  rank 0 makes the 128-byte rendezvous blob; the host hands it to the other
 *                      ranks through its control plane (the task document every worker already
 *                      polls, task.lua:96-116).
 *   mr_nccl_init       collective over cfg.world_size contexts (one per process and GPU);
 *                      cfg.rank is the NCCL rank.  Once per context.
 *   mr_shuffle_nccl    after mr_map_finish: pack -> grouped ncclSend/ncclRecv of world_size slots
 *                      -> import, all on the context's stream.  slot_bytes = 0 sizes the slots from
 *                      this job (largest segment of any rank + 25 %, one all-reduce and one host
 *                      synchronisation) and reports the size through *slot_used (may be NULL) so
 *                      that later jobs of the same shape pass it back and run with no host
 *                      synchronisation at all; overflow then surfaces as MR_ERR_OVERFLOW from
 *                      mr_job_sync, as with mr_shuffle_pack_fixed. */
#define MR_NCCL_ID_BYTES 128
int mr_nccl_unique_id(uint8_t *id /* [MR_NCCL_ID_BYTES] */);
int mr_nccl_init(mr_ctx *ctx, const uint8_t *id /* [MR_NCCL_ID_BYTES] */);
int mr_shuffle_nccl(mr_ctx *ctx, uint64_t slot_bytes, uint64_t *slot_used);

/* Peer-to-peer shuffle for ranks of one NVLink domain: no staging buffer, no collective call.  Every
 * rank owns a receive window in its HBM; the other ranks map it (CUDA IPC between processes, peer
 * access between contexts of one process) and their pack kernel stores the records straight into
 * the owner's slots over NVLink, then raises a flag there; the owner waits on its flags inside the
 * kernel that indexes the window, and sorts/reduces the records where they lie.  Same seam as above
 * (job.lua:208-221 write, job.lua:253-260 read).
 *   mr_p2p_setup   allocates this rank's window for slots of slot_bytes (multiple of 32; two
 *                  buffers of world_size slots, alternating per job) and fills *mine.
 *   (the host exchanges the 128-byte mr_peer_info blobs: every rank needs all of them, in rank order)
 *   mr_p2p_attach  maps every rank's window.  Collective in effect: call it on every rank after
 *                  every rank's mr_p2p_setup, and call neither while a shuffle is in flight.
 *   mr_shuffle_p2p after mr_map_finish, instead of the other shuffle calls; asynchronous, no host
 *                  synchronisation.  A segment larger than the slot surfaces as MR_ERR_OVERFLOW from
 *                  mr_job_sync on the sender and the owner (set the windows up again with larger
 *                  slots); a rank that never delivers surfaces as MR_ERR_CUDA after 20 s.
 * Fails with MR_ERR_UNSUPPORTED where the devices cannot map each other's memory; mr_shuffle_nccl
 * is the transport for that case.  One rank per GPU is assumed: the owner's kernel waits on the GPU
 * for its peers' records, so ranks that share one GPU (tests do) must all have finished their map side
 * (device idle) before the first of them calls mr_shuffle_p2p, or the waiting kernel can keep a peer's
 * map kernel from ever starting (reported after 20 s, not a hang). */
typedef struct mr_peer_info {
    uint64_t pid;           /* process that owns the window */
    uint64_t dev_ptr;       /* its address there (used by contexts of the same process) */
    uint64_t win_bytes;
    uint64_t slot_bytes;
    int32_t device;
    uint32_t rank;
    uint8_t ipc_handle[64]; /* cudaIpcMemHandle_t */
    uint8_t reserved[24];
} mr_peer_info;
int mr_p2p_setup(mr_ctx *ctx, uint64_t slot_bytes, mr_peer_info *mine);
int mr_p2p_attach(mr_ctx *ctx, const mr_peer_info *all /* [world_size] */);
int mr_shuffle_p2p(mr_ctx *ctx);

/* Which rank reduces a key.  Default: partition mod world_size -- every result.P<p> lives on one
 * rank, as one reduce job owns one partition in the reference (server.lua:316-324); with 10
 * partitions on 8 ranks two ranks then carry twice the others' share.  mr_shuffle_set_owners
 * replaces that by world_size-1 splitters (partition, sort word), ascending: rank r reduces the
 * keys whose (partition, sort word) lies in [splitter r-1, splitter r), i.e. a contiguous range of
 * the job's output order "partition ascending, then key ascending".  result.P<p> is then the
 * concatenation, in rank order, of the slices mr_reduce_partition(p) returns on each rank (most of
 * them empty).  The sort word of a key is its first 7 bytes, big-endian, in the top 56 bits and
 * min(length, 8) in the low byte; keys equal in it stay together.  Every rank must install the
 * same splitters (n = 0 returns to the default), between jobs.
 *   mr_row_sort_word  (partition, sort word) of row `row` of the finished job on this rank: the
 *                     host picks splitters at equal shares of a first job's rows
 *                     (lua-mapreduce-1_b200/parallel.py: balance_owners).
 *   mr_owner_of       the owner rank of a key under the installed splitters, on the host. */
int mr_shuffle_set_owners(mr_ctx *ctx, uint32_t n, const uint32_t *part, const uint64_t *word);
int mr_row_sort_word(mr_ctx *ctx, uint64_t row, uint32_t *part, uint64_t *word);
int mr_owner_of(const mr_ctx *ctx, const void *key, uint64_t klen, uint32_t *owner);
/* Reserve room for `bytes` incoming bytes; returns the device pointer to receive into. */
int mr_shuffle_recv_buffer(mr_ctx *ctx, uint64_t bytes, void **dev_ptr);
/* recv_off: host array [world_size+1] of byte offsets per source inside the receive
 * buffer; recv_nrec: host array [world_size] of records per source (the sender's
 * send_nrec for this rank, exchanged with the sizes). */
int mr_shuffle_import(mr_ctx *ctx, const uint64_t *recv_off, const uint64_t *recv_nrec);

int mr_stats_get(mr_ctx *ctx, mr_stats *out);

/* Partition id of one key on the host, same arithmetic as the device code
 * (partitionfn.lua:8-16 in IEEE binary64).  For glue-side checks. */
uint32_t mr_partition_of(const void *key, uint64_t klen, uint32_t num_reducers);

/* No counterpart in the reference (it has no device memory to manage).  Counts the
 * whitespace-separated tokens of a device buffer WITHOUT allocating anything: for hosts that
 * have given all device memory to their own pool, or that cannot take the device-wide
 * synchronisation of cudaMalloc/cudaFree where they call this.  The caller owns input and
 * output; the hash table the tokens are combined in is a fixed array inside the library
 * (4 MiB per GPU, held until exit), hence the limits: n_bytes <= MR_STATIC_MAX_BYTES, tokens
 * of at most 15 bytes, out_capacity >= number of distinct tokens (anything beyond is
 * refused with an error code and, if err_text is given, a message).
 * dev_text, dev_out, dev_n_out are device pointers; records come out in no particular
 * order; *n_tokens (host, may be NULL) receives the token count.  The call clears the shared
 * array, runs on `stream` and returns after synchronising it.  All calls on a device share
 * the array: calls whose GPU work could overlap must be serialised by the caller.  The first
 * call on each device looks up the array's address; make it before any stream capture.  That
 * first call is also when the CUDA runtime may load this library's device code onto the GPU,
 * fixed array included (lazy module loading): a host that is about to hand the rest of device
 * memory to its pool makes one call (n_bytes = 0 is enough) before it does so.
 * When out_capacity is too small the call fails with MR_ERR_NOMEM, dev_out holds out_capacity
 * valid records and *dev_n_out the number of distinct tokens found -- the capacity a retry needs. */
#define MR_STATIC_MAX_BYTES (128u * 1024u)
typedef struct mr_token_count {
    uint8_t key[15];   /* token bytes, zero padded */
    uint8_t len;       /* 1..15 */
    uint64_t count;
} mr_token_count;
int mr_count_tokens_static(const void *dev_text, uint64_t n_bytes, mr_token_count *dev_out,
                           uint64_t out_capacity, uint64_t *dev_n_out, uint64_t *n_tokens, void *stream,
                           char *err_text, uint64_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* MRGPU_H */
