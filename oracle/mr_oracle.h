/* mr_oracle.h -- CPU oracle for the word-count data plane (test infrastructure only;
 * see the header of mr_oracle.c). */
#ifndef MR_ORACLE_H
#define MR_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Rows in the reference's final order: partition ascending, key ascending. */
typedef struct mro_result {
    uint64_t n;             /* distinct keys */
    uint64_t tokens;        /* emits seen by the map side */
    uint32_t *part;         /* [n] partition id */
    uint64_t *count;        /* [n] */
    uint64_t *key_off;      /* [n+1] offsets into key_bytes */
    uint8_t *key_bytes;
    uint64_t key_bytes_len;
} mro_result;

int mro_is_space(unsigned char c);
uint32_t mro_fnv_lua52(const uint8_t *key, uint64_t len);
uint32_t mro_fnv_lua52_int(const uint8_t *key, uint64_t len);
uint32_t mro_partition(const uint8_t *key, uint64_t len, uint32_t num_reducers);
int mro_key_cmp(const uint8_t *a, uint64_t la, const uint8_t *b, uint64_t lb);

mro_result *mro_run_naive(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                          uint32_t num_reducers);
mro_result *mro_run_staged(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                           uint32_t num_reducers, int nthreads);
mro_result *mro_run_bigram(const uint8_t *const *splits, const uint64_t *lens, uint32_t nsplits,
                           uint32_t num_reducers);
mro_result *mro_groupby_u32(const uint32_t *const *keys, const uint32_t *const *vals, const uint64_t *lens,
                            uint32_t nsplits, uint32_t num_reducers, int op);
void mro_result_free(mro_result *r);

void mro_heap_sort_u64(uint64_t *v, uint32_t n);
uint32_t mro_merge_kat(const uint64_t *const *run_keys, const uint64_t *const *run_counts,
                       const uint32_t *run_lens, uint32_t nruns, uint64_t *out_key,
                       uint64_t *out_count);
int mro_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
