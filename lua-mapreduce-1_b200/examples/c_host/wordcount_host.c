/* wordcount_host.c -- a word-count worker written against include/mrgpu.h only: no Python, no torch.
 *
 * It stands where the reference's worker process stands (mapreduce/worker.lua:49-113 running the
 * job of mapreduce/job.lua:166-307 for examples/WordCount): one process per GPU, rank r of W.
 *   - the split list is the command line (taskfn.lua:3-9 emits one map job per file); rank r maps
 *     the files i with i % W == r,
 *   - the rendezvous blob of the library's own NCCL transport travels through a file, the stand-in
 *     for the task document every reference worker polls (task.lua:96-116),
 *   - each owned partition p (p % W == r) is written to <outdir>/result.P<p> as "count\tkey\n", the
 *     line format of the example's finalfn (examples/WordCount/finalfn.lua:3-9, README.md:73-76).
 *
 * usage: wordcount_host <rank> <world> <num_reducers> <idfile> <outdir> <file>...
 * Build: gcc -O2 -I include wordcount_host.c -L lua-mapreduce-1_b200 -lmrgpu -Wl,-rpath,'$ORIGIN/../..'
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "mrgpu.h"

static void die(mr_ctx *ctx, const char *what)
{
    fprintf(stderr, "wordcount_host: %s: %s\n", what, mr_last_error(ctx));
    exit(1);
}

static unsigned char *slurp(const char *path, size_t *n)
{
    FILE *f = fopen(path, "rb");
    if (!f) { perror(path); exit(1); }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char *buf = malloc(sz > 0 ? (size_t)sz : 1);
    if (sz > 0 && fread(buf, 1, (size_t)sz, f) != (size_t)sz) { perror(path); exit(1); }
    fclose(f);
    *n = (size_t)sz;
    return buf;
}

/* rank 0 publishes the blob (write to a temporary name, then rename: readers never see half of it) */
static void exchange_id(int rank, const char *idfile, uint8_t *id)
{
    if (rank == 0) {
        char tmp[4096];
        if (mr_nccl_unique_id(id)) die(NULL, "mr_nccl_unique_id");
        snprintf(tmp, sizeof tmp, "%s.tmp", idfile);
        FILE *f = fopen(tmp, "wb");
        if (!f || fwrite(id, 1, MR_NCCL_ID_BYTES, f) != MR_NCCL_ID_BYTES) { perror(tmp); exit(1); }
        fclose(f);
        if (rename(tmp, idfile)) { perror(idfile); exit(1); }
        return;
    }
    for (int tries = 0; tries < 3000; ++tries) { /* the reference's workers poll too (worker.lua:104-109) */
        FILE *f = fopen(idfile, "rb");
        if (f) {
            size_t got = fread(id, 1, MR_NCCL_ID_BYTES, f);
            fclose(f);
            if (got == MR_NCCL_ID_BYTES) return;
        }
        usleep(10000);
    }
    fprintf(stderr, "wordcount_host: no rendezvous blob at %s\n", idfile);
    exit(1);
}

int main(int argc, char **argv)
{
    if (argc < 7) {
        fprintf(stderr, "usage: %s <rank> <world> <num_reducers> <idfile> <outdir> <file>...\n", argv[0]);
        return 2;
    }
    const int rank = atoi(argv[1]), world = atoi(argv[2]);
    const unsigned R = (unsigned)atoi(argv[3]);
    const char *idfile = argv[4], *outdir = argv[5];
    char **files = argv + 6;
    const int nfiles = argc - 6;

    /* read this rank's splits first: the arena is sized from them */
    unsigned char **data = calloc((size_t)nfiles, sizeof *data);
    size_t *len = calloc((size_t)nfiles, sizeof *len);
    uint64_t total = 0;
    for (int i = rank; i < nfiles; i += world) {
        data[i] = slurp(files[i], &len[i]);
        total += len[i];
    }

    mr_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.struct_size = sizeof cfg;
    cfg.device = rank; /* one GPU per rank on one node */
    cfg.num_reducers = R;
    cfg.map_kind = MR_MAP_TOKENIZE_WS;
    cfg.reduce_kind = MR_REDUCE_SUM_U64;
    cfg.partition_kind = MR_PARTITION_FNV1_LUA52;
    cfg.arena_bytes = total + 4096u * (uint64_t)nfiles + (1u << 20);
    cfg.max_keys = 1u << 20;
    cfg.world_size = (uint32_t)world;
    cfg.rank = (uint32_t)rank;
    mr_ctx *ctx = NULL;
    if (mr_ctx_create(&cfg, &ctx)) die(NULL, "mr_ctx_create");

    if (world > 1) {
        uint8_t id[MR_NCCL_ID_BYTES];
        exchange_id(rank, idfile, id);
        if (mr_nccl_init(ctx, id)) die(ctx, "mr_nccl_init");
    }

    if (mr_job_begin(ctx, 0)) die(ctx, "mr_job_begin");
    for (int i = rank; i < nfiles; i += world)
        if (mr_map_split(ctx, (uint64_t)i, data[i], len[i])) die(ctx, "mr_map_split");
    if (mr_map_finish(ctx)) die(ctx, "mr_map_finish");
    if (world > 1 && mr_shuffle_nccl(ctx, 0, NULL)) die(ctx, "mr_shuffle_nccl");
    if (mr_reduce_all(ctx)) die(ctx, "mr_reduce_all");
    mr_job_info info;
    if (mr_job_sync(ctx, &info)) die(ctx, "mr_job_sync");

    for (unsigned p = (unsigned)rank; p < R; p += (unsigned)world) {
        mr_result *res = NULL;
        if (mr_reduce_partition(ctx, p, &res)) die(ctx, "mr_reduce_partition");
        char path[4096];
        snprintf(path, sizeof path, "%s/result.P%u", outdir, p);
        FILE *out = NULL; /* like the reference, no file for a partition that got no key (server.lua:297-324) */
        const char *key;
        uint64_t klen, count;
        while (mr_result_next(res, &key, &klen, &count) == MR_OK) {
            if (!out && !(out = fopen(path, "wb"))) { perror(path); return 1; }
            fprintf(out, "%llu\t", (unsigned long long)count);
            fwrite(key, 1, klen, out);
            fputc('\n', out);
        }
        if (out) fclose(out);
        mr_result_release(res);
    }
    printf("rank %d/%d: %llu tokens mapped, %llu keys owned\n", rank, world,
           (unsigned long long)info.n_tokens, (unsigned long long)info.n_keys);
    mr_ctx_destroy(ctx);
    return 0;
}
