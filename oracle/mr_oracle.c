/*
 * mr_oracle.c -- CPU restatement of lua-mapreduce's word-count data plane.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline.
 * The CUDA library (lua-mapreduce-1_b200/csrc) never links or calls it.
 *
 * The reference is 100% Lua and cannot run in this image (no Lua VM, no MongoDB),
 * so this file restates the algorithm from the sources, each function citing the
 * reference file:line it follows (paths relative to /root/reference).  The
 * arithmetic that lives in the third-party Lua 5.2 VM (number = IEEE double,
 * `a % b` = a - floor(a/b)*b, bit32.bxor, string `<` = strcoll in the C locale,
 * pattern class %s = isspace in the C locale) is restated from the Lua 5.2
 * manual.  Parity pin: the known answers of SURVEY.md 8c, the merge KAT of
 * mapreduce/utils.lua:360-380, the heap KAT of mapreduce/heap.lua:99-118 and
 * the keys_sorted KAT of mapreduce/utils.lua:349 (see tests/test_oracle.py).
 *
 * Two whole-job shapes, which must agree:
 *   mro_run_naive   one global table            = misc/naive.lua:1-7
 *   mro_run_staged  per-split map -> per-(partition, split) sorted runs ->
 *                   k-way heap merge + sum per partition
 *                                                = mapreduce/job.lua:154-296 +
 *                                                  mapreduce/utils.lua:206-271
 * Both deliver (partition, key, count) rows ordered as the reference's final
 * iterator sees them: partitions ascending (server.lua:365), keys ascending in
 * unsigned-byte order inside a partition (utils.lua:126, utils.lua:214).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "mr_oracle.h"

/* ---------------------------------------------------------------- tokeniser */

/* examples/WordCount/mapfn.lua:4-5: io.lines + gmatch("[^%s]+").  %s is C-locale
 * isspace: 0x20 and 0x09..0x0D.  '\n' is itself whitespace, so the token stream
 * of a split is the split cut at those six bytes with empty pieces dropped. */
int mro_is_space(unsigned char c) { return c == 0x20 || (c >= 0x09 && c <= 0x0D); }

/* ---------------------------------------------------------------- partitioner */

/* examples/WordCount/partitionfn.lua:10-15 in IEEE binary64:
 *   h = offset_basis; for each byte: h = (h * FNV_prime) % 2^32; h = bxor(h, byte)
 * Lua 5.2's `%` on numbers is a - floor(a/b)*b.  The product exceeds 2^53 so it
 * rounds (to nearest even); this is therefore not uint32 FNV-1. */
uint32_t mro_fnv_lua52(const uint8_t *key, uint64_t len)
{
    volatile double h = 2166136261.0;
    const double prime = 16777619.0, max = 4294967296.0;
    for (uint64_t i = 0; i < len; ++i) {
        volatile double p = h * prime; /* volatile: one rounding, no FMA, no x87 excess */
        volatile double q = floor(p / max);
        double m = p - q * max;
        uint32_t u = (uint32_t)m; /* bit32.bxor takes the number mod 2^32 */
        h = (double)(u ^ (uint32_t)key[i]);
    }
    return (uint32_t)h;
}

/* Same function with 64-bit integers only: round the 57-bit product to 53
 * significant bits, ties to even, then keep the low 32 bits.  Used to check the
 * device code's arithmetic independently of the host FPU. */
uint32_t mro_fnv_lua52_int(const uint8_t *key, uint64_t len)
{
    uint64_t h = 2166136261u;
    for (uint64_t i = 0; i < len; ++i) {
        uint64_t p = h * 16777619u;
        int top = 63 - __builtin_clzll(p | 1);
        if (top >= 53) {
            int s = top - 52;
            uint64_t half = 1ull << (s - 1), rem = p & ((1ull << s) - 1);
            p >>= s;
            if (rem > half || (rem == half && (p & 1))) p += 1;
            p <<= s;
        }
        h = (uint32_t)p ^ key[i];
    }
    return (uint32_t)h;
}

/* partitionfn.lua:15 `h % NUM_REDUCERS` with Lua's float modulo; job.lua:203-207
 * then requires an integer-valued number. */
uint32_t mro_partition(const uint8_t *key, uint64_t len, uint32_t num_reducers)
{
    double h = (double)mro_fnv_lua52(key, len), r = (double)num_reducers;
    double m = h - floor(h / r) * r;
    return (uint32_t)m;
}

/* ---------------------------------------------------------------- key order */

/* Lua string `<` (utils.lua:126 table.sort, utils.lua:214 heap cmp) is strcoll per
 * NUL-free segment; standalone lua never calls setlocale, so the C locale applies:
 * unsigned-byte lexicographic, the shorter string first on an equal prefix. */
int mro_key_cmp(const uint8_t *a, uint64_t la, const uint8_t *b, uint64_t lb)
{
    uint64_t n = la < lb ? la : lb;
    int c = n ? memcmp(a, b, n) : 0;
    if (c) return c;
    return la < lb ? -1 : (la > lb ? 1 : 0);
}

/* ---------------------------------------------------------------- key table */

/* The Lua table `result[key]` of job.lua:83-97 (and `vocab[w]` of naive.lua:4):
 * an open-addressing map from byte string to a running sum.  The value list the
 * reference keeps per key is summed by the combiner (reducefn.lua:1-5) either
 * inline past MAX_MAP_RESULT (job.lua:92-96) or at job.lua:198-202; for a sum the
 * grouping does not change the result, so the list is kept as its sum. */
typedef struct {
    const uint8_t *ptr;
    uint32_t len;
    uint32_t hash;
    uint64_t count;
} kt_slot;

typedef struct {
    kt_slot *slots;
    uint64_t cap, used;
} ktable;

static uint32_t kt_hash(const uint8_t *p, uint64_t n)
{
    uint64_t h = 0x9E3779B97F4A7C15ull ^ n;
    for (uint64_t i = 0; i < n; ++i) h = (h ^ p[i]) * 0x100000001B3ull;
    h ^= h >> 29;
    return (uint32_t)(h ^ (h >> 32)) | 1u;
}

static void kt_init(ktable *t, uint64_t cap)
{
    t->cap = cap;
    t->used = 0;
    t->slots = (kt_slot *)calloc(cap, sizeof(kt_slot));
}

static void kt_add(ktable *t, const uint8_t *p, uint32_t n, uint64_t c);

static void kt_grow(ktable *t)
{
    ktable big;
    kt_init(&big, t->cap * 2);
    for (uint64_t i = 0; i < t->cap; ++i)
        if (t->slots[i].hash) kt_add(&big, t->slots[i].ptr, t->slots[i].len, t->slots[i].count);
    free(t->slots);
    *t = big;
}

static void kt_add(ktable *t, const uint8_t *p, uint32_t n, uint64_t c)
{
    if ((t->used + 1) * 2 > t->cap) kt_grow(t);
    uint32_t h = kt_hash(p, n);
    uint64_t i = h & (t->cap - 1);
    for (;;) {
        kt_slot *s = &t->slots[i];
        if (!s->hash) {
            s->ptr = p; s->len = n; s->hash = h; s->count = c;
            t->used++;
            return;
        }
        if (s->hash == h && s->len == n && memcmp(s->ptr, p, n) == 0) {
            s->count += c;
            return;
        }
        i = (i + 1) & (t->cap - 1);
    }
}

/* mapfn.lua:3-9 feeding emit (job.lua:83-97): every maximal run of non-space
 * bytes of one split is emit(w, 1).  Tokens never span splits. */
static uint64_t map_split(ktable *t, const uint8_t *b, uint64_t n)
{
    uint64_t i = 0, tokens = 0;
    while (i < n) {
        while (i < n && mro_is_space(b[i])) ++i;
        uint64_t s = i;
        while (i < n && !mro_is_space(b[i])) ++i;
        if (i > s) { kt_add(t, b + s, (uint32_t)(i - s), 1); ++tokens; }
    }
    return tokens;
}

/* ---------------------------------------------------------------- rows */

typedef struct {
    const uint8_t *ptr;
    uint32_t len;
    uint32_t part;
    uint64_t count;
} row;

static int row_cmp_key(const void *x, const void *y)
{
    const row *a = (const row *)x, *b = (const row *)y;
    return mro_key_cmp(a->ptr, a->len, b->ptr, b->len);
}

static int row_cmp_part_key(const void *x, const void *y)
{
    const row *a = (const row *)x, *b = (const row *)y;
    if (a->part != b->part) return a->part < b->part ? -1 : 1;
    return mro_key_cmp(a->ptr, a->len, b->ptr, b->len);
}

static mro_result *result_from_rows(const row *rows, uint64_t n, uint64_t tokens)
{
    mro_result *r = (mro_result *)calloc(1, sizeof(*r));
    r->n = n;
    r->tokens = tokens;
    r->part = (uint32_t *)malloc((n + 1) * sizeof(uint32_t));
    r->count = (uint64_t *)malloc((n + 1) * sizeof(uint64_t));
    r->key_off = (uint64_t *)malloc((n + 1) * sizeof(uint64_t));
    uint64_t total = 0;
    for (uint64_t i = 0; i < n; ++i) total += rows[i].len;
    r->key_bytes_len = total;
    r->key_bytes = (uint8_t *)malloc(total + 1);
    uint64_t off = 0;
    for (uint64_t i = 0; i < n; ++i) {
        r->part[i] = rows[i].part;
        r->count[i] = rows[i].count;
        r->key_off[i] = off;
        memcpy(r->key_bytes + off, rows[i].ptr, rows[i].len);
        off += rows[i].len;
    }
    r->key_off[n] = off;
    return r;
}

void mro_result_free(mro_result *r)
{
    if (!r) return;
    free(r->part); free(r->count); free(r->key_off); free(r->key_bytes); free(r);
}

static row *rows_from_table(const ktable *t, uint32_t num_reducers)
{
    row *rows = (row *)malloc((t->used + 1) * sizeof(row));
    uint64_t k = 0;
    for (uint64_t i = 0; i < t->cap; ++i) {
        const kt_slot *s = &t->slots[i];
        if (!s->hash) continue;
        rows[k].ptr = s->ptr; rows[k].len = s->len; rows[k].count = s->count;
        rows[k].part = mro_partition(s->ptr, s->len, num_reducers);
        ++k;
    }
    return rows;
}

/* ---------------------------------------------------------------- naive */

/* misc/naive.lua:1-7: one table over the whole input; the reference's own
 * correctness oracle in test.sh:11-15.  Rows are then put in the final
 * iterator's order so they can be compared with the staged path. */
mro_result *mro_run_naive(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                          uint32_t num_reducers)
{
    ktable t;
    kt_init(&t, 1u << 16);
    uint64_t tokens = 0;
    for (uint32_t s = 0; s < nsplits; ++s) tokens += map_split(&t, splits[s], lens[s]);
    row *rows = rows_from_table(&t, num_reducers);
    qsort(rows, t.used, sizeof(row), row_cmp_part_key);
    mro_result *r = result_from_rows(rows, t.used, tokens);
    free(rows);
    free(t.slots);
    return r;
}

/* ---------------------------------------------------------------- heap */

/* mapreduce/heap.lua:29-70: binary min-heap in a 1-based array; push bubbles the
 * new item up while it is smaller than its parent; pop moves the last item to the
 * root and sinks it, preferring the right child only when it is strictly smaller
 * than the left.  The comparison is utils.lua:214 (key `<`). */
typedef struct {
    const row *r;   /* current head row of the run */
    uint32_t which; /* run index */
} hitem;

typedef struct {
    hitem *d; /* d[1..n] */
    uint32_t n;
} heap;

static int h_less(const hitem *a, const hitem *b)
{
    return mro_key_cmp(a->r->ptr, a->r->len, b->r->ptr, b->r->len) < 0;
}

static void heap_push(heap *h, hitem v)
{
    uint32_t pos = ++h->n;
    while (pos > 1) {
        uint32_t p = pos / 2;
        if (!h_less(&v, &h->d[p])) break;
        h->d[pos] = h->d[p];
        pos = p;
    }
    h->d[pos] = v;
}

static void heap_pop(heap *h)
{
    hitem v = h->d[h->n--];
    if (h->n == 0) return;
    uint32_t pos = 1;
    for (;;) {
        uint32_t l = 2 * pos, r = l + 1;
        if (l > h->n) break;
        uint32_t child = (r <= h->n && h_less(&h->d[r], &h->d[l])) ? r : l;
        if (!h_less(&h->d[child], &v)) break;
        h->d[pos] = h->d[child];
        pos = child;
    }
    h->d[pos] = v;
}

/* Exposed for the KAT of heap.lua:99-118 (push 20,10,15,1 -> pops ascending). */
void mro_heap_sort_u64(uint64_t *v, uint32_t n)
{
    /* keys are compared as 8-byte big-endian strings = numeric order */
    row *rows = (row *)malloc((n + 1) * sizeof(row));
    uint8_t *buf = (uint8_t *)malloc((size_t)n * 8 + 8);
    heap h = { (hitem *)malloc((n + 2) * sizeof(hitem)), 0 };
    for (uint32_t i = 0; i < n; ++i) {
        for (int b = 0; b < 8; ++b) buf[i * 8 + b] = (uint8_t)(v[i] >> (56 - 8 * b));
        rows[i].ptr = buf + i * 8; rows[i].len = 8; rows[i].count = v[i];
        hitem it = { &rows[i], i };
        heap_push(&h, it);
    }
    for (uint32_t i = 0; i < n; ++i) { v[i] = h.d[1].r->count; heap_pop(&h); }
    free(h.d); free(buf); free(rows);
}

/* ---------------------------------------------------------------- merge */

typedef struct {
    row *rows;
    uint64_t n;
} run;

/* utils.lua:206-271 merge_iterator + job.lua:264-284 reduce loop over the runs of
 * ONE partition: a heap holds the head of every run (take_next, 218-230);
 * merge_min_keys (232-247) pops the minimum and concatenates the values of every
 * other head with the same key, refilling from the run each came from; the
 * reducer (reducefn.lua:1-5) sums them.  The assoc/commut/idempotent short-cut
 * (job.lua:266) skips the call for a single value, which for a sum is identity.
 * Appends to out[], returns the new length. */
static uint64_t merge_partition(const run *runs, uint32_t nruns, row *out, uint64_t nout)
{
    heap h = { (hitem *)malloc((nruns + 2) * sizeof(hitem)), 0 };
    uint64_t *pos = (uint64_t *)calloc(nruns + 1, sizeof(uint64_t));
    for (uint32_t i = 0; i < nruns; ++i)
        if (runs[i].n) { hitem it = { &runs[i].rows[0], i }; pos[i] = 1; heap_push(&h, it); }
    while (h.n) {
        hitem top = h.d[1];
        heap_pop(&h);
        row acc = *top.r;
        if (pos[top.which] < runs[top.which].n) {
            hitem it = { &runs[top.which].rows[pos[top.which]++], top.which };
            heap_push(&h, it);
        }
        while (h.n && mro_key_cmp(acc.ptr, acc.len, h.d[1].r->ptr, h.d[1].r->len) == 0) {
            hitem aux = h.d[1];
            heap_pop(&h);
            acc.count += aux.r->count;
            if (pos[aux.which] < runs[aux.which].n) {
                hitem it = { &runs[aux.which].rows[pos[aux.which]++], aux.which };
                heap_push(&h, it);
            }
        }
        out[nout++] = acc;
    }
    free(pos);
    free(h.d);
    return nout;
}

/* KAT entry for utils.lua:360-380: numeric keys (Lua compares numbers numerically),
 * each run given as parallel (key, number-of-values) arrays; the merged value
 * lists are all 1s in that vector, so their length is their sum.  Returns the
 * number of merged keys; out_key/out_count receive them in merge order. */
uint32_t mro_merge_kat(const uint64_t *const *run_keys, const uint64_t *const *run_counts,
                       const uint32_t *run_lens, uint32_t nruns, uint64_t *out_key,
                       uint64_t *out_count)
{
    run *runs = (run *)calloc(nruns + 1, sizeof(run));
    uint64_t total = 0;
    for (uint32_t i = 0; i < nruns; ++i) total += run_lens[i];
    uint8_t *buf = (uint8_t *)malloc(total * 8 + 8);
    uint64_t o = 0;
    for (uint32_t i = 0; i < nruns; ++i) {
        runs[i].n = run_lens[i];
        runs[i].rows = (row *)malloc((run_lens[i] + 1) * sizeof(row));
        for (uint32_t j = 0; j < run_lens[i]; ++j, ++o) {
            for (int b = 0; b < 8; ++b) buf[o * 8 + b] = (uint8_t)(run_keys[i][j] >> (56 - 8 * b));
            runs[i].rows[j].ptr = buf + o * 8; runs[i].rows[j].len = 8;
            runs[i].rows[j].count = run_counts[i][j]; runs[i].rows[j].part = 0;
        }
    }
    row *out = (row *)malloc((total + 1) * sizeof(row));
    uint64_t n = merge_partition(runs, nruns, out, 0);
    for (uint64_t i = 0; i < n; ++i) {
        uint64_t k = 0;
        for (int b = 0; b < 8; ++b) k = (k << 8) | out[i].ptr[b];
        out_key[i] = k; out_count[i] = out[i].count;
    }
    for (uint32_t i = 0; i < nruns; ++i) free(runs[i].rows);
    free(out); free(buf); free(runs);
    return (uint32_t)n;
}

/* ---------------------------------------------------------------- staged */

/* The reference's real shape.
 * Map job per split (job.lua:166-227): tokenise into a per-job table, sort its
 * keys (keys_sorted, utils.lua:123-128), combine, partition each key
 * (job.lua:203-207) and append it to the run P<part>.M<split> (208-214).
 * server_prepare_reduce (server.lua:279-329) makes one reduce job per partition
 * that has at least one run; partitions nobody wrote do not exist.
 * Reduce job per partition (job.lua:232-295): merge_partition above.
 * server_final (server.lua:346-385) walks result files sorted by zero-padded
 * partition index.  Threads stand in for the reference's worker processes. */
mro_result *mro_run_staged(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                           uint32_t num_reducers, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    uint32_t R = num_reducers;
    run *runs = (run *)calloc((size_t)nsplits * R + 1, sizeof(run)); /* [split][part] */
    row **split_rows = (row **)calloc(nsplits + 1, sizeof(row *));
    uint64_t tokens = 0;

#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) reduction(+ : tokens)
    for (uint32_t s = 0; s < nsplits; ++s) {
        ktable t;
        kt_init(&t, 1u << 14);
        tokens += map_split(&t, splits[s], lens[s]);
        row *rows = rows_from_table(&t, R);
        qsort(rows, t.used, sizeof(row), row_cmp_key);
        /* stable split of the key-sorted list into per-partition runs */
        uint64_t *cnt = (uint64_t *)calloc(R + 1, sizeof(uint64_t));
        for (uint64_t i = 0; i < t.used; ++i) cnt[rows[i].part]++;
        row *byp = (row *)malloc((t.used + 1) * sizeof(row));
        uint64_t off = 0;
        for (uint32_t p = 0; p < R; ++p) {
            runs[(size_t)s * R + p].rows = byp + off;
            runs[(size_t)s * R + p].n = 0;
            off += cnt[p];
        }
        for (uint64_t i = 0; i < t.used; ++i) {
            run *rn = &runs[(size_t)s * R + rows[i].part];
            rn->rows[rn->n++] = rows[i];
        }
        split_rows[s] = byp;
        free(cnt); free(rows); free(t.slots);
    }

    /* reduce: one job per partition */
    uint64_t *part_total = (uint64_t *)calloc(R + 1, sizeof(uint64_t));
    for (uint32_t s = 0; s < nsplits; ++s)
        for (uint32_t p = 0; p < R; ++p) part_total[p] += runs[(size_t)s * R + p].n;
    row **part_out = (row **)calloc(R + 1, sizeof(row *));
    uint64_t *part_n = (uint64_t *)calloc(R + 1, sizeof(uint64_t));

#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
    for (uint32_t p = 0; p < R; ++p) {
        if (!part_total[p]) continue;
        run *mine = (run *)malloc((nsplits + 1) * sizeof(run));
        for (uint32_t s = 0; s < nsplits; ++s) mine[s] = runs[(size_t)s * R + p];
        part_out[p] = (row *)malloc((part_total[p] + 1) * sizeof(row));
        part_n[p] = merge_partition(mine, nsplits, part_out[p], 0);
        free(mine);
    }

    uint64_t n = 0;
    for (uint32_t p = 0; p < R; ++p) n += part_n[p];
    row *all = (row *)malloc((n + 1) * sizeof(row));
    uint64_t k = 0;
    for (uint32_t p = 0; p < R; ++p) {
        if (part_n[p]) memcpy(all + k, part_out[p], part_n[p] * sizeof(row));
        k += part_n[p];
        free(part_out[p]);
    }
    mro_result *r = result_from_rows(all, n, tokens);
    free(all); free(part_out); free(part_n); free(part_total);
    for (uint32_t s = 0; s < nsplits; ++s) free(split_rows[s]);
    free(split_rows); free(runs);
    return r;
}

/* ---------------------------------------------------------------- bigram count */

/* BASELINE config 4: the WordCount job with a mapfn that emits word pairs,
 *     for line in io.lines(value) do
 *       local prev
 *       for w in line:gmatch("[^%s]+") do if prev then emit(prev .. " " .. w, 1) end; prev = w end
 *     end
 * i.e. mapfn.lua:4-8 with the key `w_i .. " " .. w_{i+1}` inside one line (io.lines cuts at '\n'
 * only; '\r' stays in the line and is whitespace for %s).  Everything downstream is the WordCount
 * path: emit + sum (job.lua:83-97, reducefn.lua:1-5), partitionfn.lua:8-16, Lua string order. */
mro_result *mro_run_bigram(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                           uint32_t num_reducers)
{
    ktable t;
    kt_init(&t, 1u << 16);
    uint64_t pairs = 0, cap = 1u << 20, used = 0, nblocks = 0, blocks_cap = 64;
    uint8_t **blocks = (uint8_t **)malloc(blocks_cap * sizeof(uint8_t *)); /* key bytes live in blocks */
    uint8_t *cur = (uint8_t *)malloc(cap);
    blocks[nblocks++] = cur;
    for (uint32_t s = 0; s < nsplits; ++s) {
        const uint8_t *b = splits[s];
        uint64_t n = lens[s], i = 0;
        const uint8_t *prev = NULL;
        uint64_t prev_len = 0;
        while (i < n) {
            while (i < n && mro_is_space(b[i])) { if (b[i] == '\n') prev = NULL; ++i; }
            uint64_t st = i;
            while (i < n && !mro_is_space(b[i])) ++i;
            if (i == st) break;
            if (prev) {
                uint64_t klen = prev_len + 1 + (i - st);
                if (used + klen > cap) {
                    if (nblocks == blocks_cap) { blocks_cap *= 2; blocks = (uint8_t **)realloc(blocks, blocks_cap * sizeof(uint8_t *)); }
                    cap = klen > (1u << 20) ? klen : (1u << 20);
                    cur = (uint8_t *)malloc(cap); used = 0; blocks[nblocks++] = cur;
                }
                uint8_t *k = cur + used;
                memcpy(k, prev, prev_len); k[prev_len] = ' '; memcpy(k + prev_len + 1, b + st, i - st);
                uint64_t before = t.used;
                kt_add(&t, k, (uint32_t)klen, 1);
                if (t.used != before) used += klen; /* keep the bytes only for a new key */
                ++pairs;
            }
            prev = b + st; prev_len = i - st;
        }
    }
    row *rows = rows_from_table(&t, num_reducers);
    qsort(rows, t.used, sizeof(row), row_cmp_part_key);
    mro_result *r = result_from_rows(rows, t.used, pairs);
    free(rows); free(t.slots);
    for (uint64_t j = 0; j < nblocks; ++j) free(blocks[j]);
    free(blocks);
    return r;
}

/* ---------------------------------------------------------------- group-by sum on number keys */

/* BASELINE config 5: a job whose mapfn emits (number key, number value) pairs, whose
 * partitionfn is `key % NUM_REDUCERS` and whose reducer is the WordCount sum
 * (reducefn.lua:1-5).  The framework treats it like any other job: emit groups values by key
 * (job.lua:83-97), keys are ordered with Lua `<`, which on numbers is numeric (utils.lua:126,
 * utils.lua:214), partitions ascending (server.lua:365).  Keys are carried as 4 big-endian bytes
 * in the rows so that byte order equals numeric order. */
typedef struct { uint32_t key; uint32_t used; uint64_t sum; } gb_slot;

static int gb_cmp(const void *x, const void *y)
{
    const row *a = (const row *)x, *b = (const row *)y;
    if (a->part != b->part) return a->part < b->part ? -1 : 1;
    return memcmp(a->ptr, b->ptr, 4);
}

/* op: 0 sum (reducefn.lua:1-5), 1 min, 2 max, 3 count of values */
mro_result *mro_groupby_u32(const uint32_t *const *keys, const uint32_t *const *vals, const uint64_t *lens,
                            uint32_t nsplits, uint32_t num_reducers, int op)
{
    uint64_t cap = 1u << 16, used = 0, records = 0;
    gb_slot *t = (gb_slot *)calloc(cap, sizeof(gb_slot));
    for (uint32_t s = 0; s < nsplits; ++s)
        for (uint64_t i = 0; i < lens[s]; ++i) {
            if ((used + 1) * 2 > cap) {
                gb_slot *big = (gb_slot *)calloc(cap * 2, sizeof(gb_slot));
                for (uint64_t j = 0; j < cap; ++j)
                    if (t[j].used) {
                        uint64_t h = (t[j].key * 0x9E3779B1u) & (cap * 2 - 1);
                        while (big[h].used) h = (h + 1) & (cap * 2 - 1);
                        big[h] = t[j];
                    }
                free(t); t = big; cap *= 2;
            }
            uint32_t k = keys[s][i];
            uint64_t h = (k * 0x9E3779B1u) & (cap - 1);
            while (t[h].used && t[h].key != k) h = (h + 1) & (cap - 1);
            uint64_t v = vals[s][i];
            if (!t[h].used) { t[h].used = 1; t[h].key = k; ++used; t[h].sum = op == 1 ? v : (op == 2 ? v : 0); }
            if (op == 0) t[h].sum += v;
            else if (op == 1) t[h].sum = v < t[h].sum ? v : t[h].sum;
            else if (op == 2) t[h].sum = v > t[h].sum ? v : t[h].sum;
            else t[h].sum += 1;
            ++records;
        }
    row *rows = (row *)malloc((used + 1) * sizeof(row));
    uint8_t *kb = (uint8_t *)malloc(used * 4 + 4);
    uint64_t n = 0;
    for (uint64_t j = 0; j < cap; ++j)
        if (t[j].used) {
            uint32_t k = t[j].key;
            kb[4 * n] = (uint8_t)(k >> 24); kb[4 * n + 1] = (uint8_t)(k >> 16);
            kb[4 * n + 2] = (uint8_t)(k >> 8); kb[4 * n + 3] = (uint8_t)k;
            rows[n].ptr = kb + 4 * n; rows[n].len = 4; rows[n].count = t[j].sum;
            double kd = (double)k, r = (double)num_reducers;   /* Lua: key % NUM_REDUCERS on doubles */
            rows[n].part = (uint32_t)(kd - floor(kd / r) * r);
            ++n;
        }
    qsort(rows, n, sizeof(row), gb_cmp);
    mro_result *res = result_from_rows(rows, n, records);
    free(rows); free(kb); free(t);
    return res;
}

int mro_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
