/*
 * corpus.c -- deterministic synthetic "Europarl-shaped" text for the word-count
 * benchmark.  Integer arithmetic only, so the bytes
 * are identical on every machine for a given (seed, split id, line count).
 *
 * Shape (reference README.md:43-45): 197 splits of <=10,000 lines, ~25 tokens per
 * line, token ranks Zipf(s=1) over 2^19 types; the word of a type is a lowercase
 * string whose length grows with the rank (stream mean ~5.0 bytes); one 0x20
 * between tokens, 0x0A at line end.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MRC_V (1u << 19)
#define MRC_MAXLEN 14

static const uint32_t class_size[MRC_MAXLEN + 1] = {
    0, 6, 24, 80, 300, 1100, 4000, 14000, 40000, 90000, 130000, 110000, 70000, 40000,
    0 /* length 14 takes the rest */
};

static uint8_t (*g_word)[16]; /* [V][16], zero padded */
static uint8_t *g_len;        /* [V] */
static uint64_t *g_cum;       /* [V] inclusive cumulative integer weights */
static uint64_t g_total;
static volatile int g_ready;

static inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int mrc_init(void)
{
    if (g_ready) return 0;
    uint8_t(*word)[16] = calloc(MRC_V, 16);
    uint8_t *len = malloc(MRC_V);
    uint64_t *cum = malloc(sizeof(uint64_t) * MRC_V);
    if (!word || !len || !cum) return -1;
    uint32_t t = 0;
    for (int L = 1; L <= MRC_MAXLEN; ++L) {
        uint32_t n = (L == MRC_MAXLEN) ? MRC_V - t : class_size[L];
        int Ls = L > 13 ? 13 : L; /* 26^13 < 2^63; longer words repeat the index */
        uint64_t mod = 1;
        for (int i = 0; i < Ls; ++i) mod *= 26;
        /* odd multiplier that is not a multiple of 13: a bijection mod 26^Ls */
        uint64_t a = (0x5DEECE66Dull * (uint64_t)(2 * L + 1)) | 1ull;
        while (a % 13 == 0) a += 2;
        a %= mod;
        if (a % 2 == 0 || a % 13 == 0) a = 1;
        for (uint32_t j = 0; j < n; ++j, ++t) {
            uint64_t x = (unsigned __int128)j * a % mod;
            x = (x + (uint64_t)L * 7919u) % mod;
            for (int i = Ls - 1; i >= 0; --i) { word[t][i] = (uint8_t)('a' + x % 26); x /= 26; }
            for (int i = Ls; i < L; ++i) word[t][i] = (uint8_t)('a' + (j + i) % 26);
            len[t] = (uint8_t)L;
        }
    }
    uint64_t acc = 0;
    for (uint32_t r = 0; r < MRC_V; ++r) { acc += (1ull << 40) / (r + 1); cum[r] = acc; }
    g_word = word; g_len = len; g_cum = cum; g_total = acc;
    __sync_synchronize();
    g_ready = 1;
    return 0;
}

uint64_t mrc_max_bytes(uint32_t lines) { return (uint64_t)lines * 40u * (MRC_MAXLEN + 1) + 16; }

/* Writes split `split_id` (lines lines) into out[0..cap); returns bytes written
 * (0 on overflow) and the token count through *tokens. */
uint64_t mrc_generate_split(uint64_t seed, uint32_t split_id, uint32_t lines, uint8_t *out,
                            uint64_t cap, uint64_t *tokens)
{
    if (mrc_init()) return 0;
    uint64_t s = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(split_id + 1));
    (void)splitmix64(&s);
    uint64_t o = 0, ntok = 0;
    for (uint32_t l = 0; l < lines; ++l) {
        uint32_t k = 10 + (uint32_t)(splitmix64(&s) % 31);
        if (o + (uint64_t)k * (MRC_MAXLEN + 1) + 1 > cap) return 0;
        for (uint32_t i = 0; i < k; ++i) {
            uint64_t u = splitmix64(&s) % g_total;
            uint32_t lo = 0, hi = MRC_V - 1; /* first r with cum[r] > u */
            while (lo < hi) {
                uint32_t mid = (lo + hi) >> 1;
                if (g_cum[mid] > u) hi = mid; else lo = mid + 1;
            }
            memcpy(out + o, g_word[lo], 16 <= cap - o ? 16 : g_len[lo]);
            o += g_len[lo];
            out[o++] = (i + 1 == k) ? '\n' : ' ';
        }
        ntok += k;
    }
    if (tokens) *tokens = ntok;
    return o;
}

/* word of a type, for tests */
uint32_t mrc_word(uint32_t type, uint8_t *out16)
{
    if (mrc_init() || type >= MRC_V) return 0;
    memcpy(out16, g_word[type], 16);
    return g_len[type];
}
