/*
 * oracle/spl_oracle.h -- TEST INFRASTRUCTURE ONLY (CPU restatement of the reference hot path).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (libsplinter_gpu.so) never links it and has
 * no CPU fallback.
 *
 * Parity status: PINNED.  Every function here is checked against the reference's own
 * code compiled unmodified into oracle/_ref/ (tests/test_oracle_pinned.py, run in the
 * build container where /root/reference exists) and against golden vectors generated
 * from that compiled reference (tests/golden/, generator tests/golden/make_golden.py).
 * The reference's own test-suite holds no golden vector for this path.
 */
#ifndef SPL_ORACLE_H
#define SPL_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- store layout (restates splinter.h:176-255; offsets measured with offsetof) ---- */
#define SPLO_MAGIC        0x534C4E54u /* splinter.h:29 */
#define SPLO_VERSION      4u          /* splinter.h:32 */
#define SPLO_HEADER_SIZE  5440u
#define SPLO_SLOT_META    128u        /* bytes before embedding[] */
#define SPLO_KEY_MAX      64u
#define SPLO_OFF_HASH     0u
#define SPLO_OFF_EPOCH    8u
#define SPLO_OFF_VAL_OFF  16u
#define SPLO_OFF_VAL_LEN  20u
#define SPLO_OFF_TYPE     24u
#define SPLO_OFF_BLOOM    56u
#define SPLO_OFF_KEY      64u
#define SPLO_OFF_EMBED    128u

/* result ordering */
#define SPLO_ORDER_COSINE 0 /* reference: similarity desc, distance asc (_search.c:64-72), then slot idx asc */
#define SPLO_ORDER_L2     1 /* restated for the batched L2 config: distance asc, similarity desc, idx asc */

/* dense-row flag bits (same meaning as the device mirror's flags byte) */
#define SPLO_F_LIST  0x01 /* hash!=0 && val_len>0 && epoch!=0 && key[0]   (splinter_list + _search.c:366,375) */
#define SPLO_F_ENUM  0x02 /* hash!=0 && epoch!=0 && key[0]                (enumerate_matches + same)          */
#define SPLO_F_EMB   0x04 /* an even-epoch embedding copy exists          (splinter_get_embedding == 0)       */

uint64_t splo_fnv1a(const char *s);                                  /* splinter.c:67-71 */
float    splo_cosine(const float *a, const float *b, int dim);       /* _search.c:43-53  */
float    splo_euclid(const float *a, const float *b, int dim);       /* _search.c:55-62  */
/* <0: a sorts before b.  Total order = reference comparator + idx ascending for ties */
int      splo_compare(int order, float sim_a, float dist_a, uint32_t idx_a,
                      float sim_b, float dist_b, uint32_t idx_b);

/* ---- store image: create / mutate with the reference's semantics (single-threaded) ---- */
size_t splo_image_size(uint32_t slots, uint32_t max_val_sz, uint32_t dim);
int    splo_image_init(void *base, uint32_t slots, uint32_t max_val_sz, uint32_t dim);   /* splinter.c:148-233 */
long   splo_image_find(const void *base, const char *key);                               /* probe, -1 if absent */
long   splo_image_set(void *base, const char *key, const void *val, uint32_t len);       /* splinter.c:365-429; returns slot */
int    splo_image_set_embedding(void *base, const char *key, const float *vec);          /* splinter.c:567-588 */
int    splo_image_set_label(void *base, const char *key, uint64_t mask);                 /* splinter.c:835-850 */
int    splo_image_unset_label(void *base, const char *key, uint64_t mask);               /* splinter.c:852-867 */
int    splo_image_retrain(void *base, const char *key);                                  /* splinter.c:799-833 */
int    splo_image_unset(void *base, const char *key);                                    /* splinter.c:326-363 */
uint32_t splo_image_dim(const void *base, size_t file_size);                             /* geometry gate, SURVEY 8(b) */

/* Bulk synthetic store of SURVEY 8(d): keys doc%08d inserted by linear probing like
 * splinter_set, embeddings/bloom generated from (seed, slot) by the counter RNG below. */
long   splo_image_fill_synthetic(void *base, uint32_t nkeys, uint64_t seed_corpus,
                                 uint64_t seed_bloom, int threads);

/* ---- candidate enumeration over an image (slot-index order) ---- */
size_t splo_enumerate_matches(const void *base, uint64_t mask, uint32_t *out_idx);       /* splinter.c:977-989 */
size_t splo_list(const void *base, uint32_t *out_idx);                                   /* splinter.c:466-477 */

/* seqlock reader rule (splinter.c:598-605): 0 = consistent copy, -1 = busy/torn */
int    splo_snapshot_row(const void *base, uint32_t idx, float *row_out, uint64_t *epoch_out);

/* image -> dense arrays exactly as the GPU mirror must hold them */
void   splo_image_to_dense(const void *base, float *E, uint64_t *bloom, uint8_t *flags, uint64_t *epoch);

/* ---- search ---- */
typedef struct {
    uint32_t idx;
    float    similarity;
    float    distance;
    uint32_t has_embedding;
} splo_result;

/* cmd_search policy (_search.c:329-416) over a store image. exclude_key may be NULL.
 * limit<=0 = all.  Returns result count written (<= limit); *survivors = rows before limit. */
long splo_search_image(const void *base, const float *q, int order, int limit, uint64_t bloom_mask,
                       float min_sim, float max_dist, const char *exclude_key,
                       splo_result *out, long *survivors);

/* same policy over dense arrays (flags as SPLO_F_*); nq queries, out[nq][limit], counts[nq] */
long splo_search_dense(const float *q, uint32_t nq, const float *E, const uint64_t *bloom,
                       const uint8_t *flags, uint64_t n, uint32_t dim, int order, int limit,
                       uint64_t bloom_mask, float min_sim, float max_dist, uint32_t exclude_idx,
                       int threads, splo_result *out, uint32_t *counts);

/* k-way merge of per-shard top-k lists (multi-GPU restatement) */
uint32_t splo_merge_topk(const splo_result *lists, const uint32_t *counts, uint32_t nlists,
                         uint32_t stride, int order, uint32_t k, splo_result *out);

/* ---- synthetic data (counter RNG; the CUDA generator reproduces these bit-for-bit) ---- */
#define SPLO_CLASS_EMPTY 0 /* 0.1 %: hash == 0                    */
#define SPLO_CLASS_ZERO  1 /* 2 %  : live key, never embedded     */
#define SPLO_CLASS_EMBED 2
uint64_t splo_mix64(uint64_t z);
int      splo_synth_class(uint64_t seed, uint64_t row);
void     splo_synth_row(uint64_t seed, uint64_t row, uint32_t dim, float *out); /* unit-norm, uniform[-1,1) */
uint64_t splo_synth_bloom(uint64_t seed, uint64_t row);   /* bit b: p = 2^-(1 + b%4), b<63; bit 63 always */
void     splo_synth_dense(uint64_t seed_corpus, uint64_t seed_bloom, uint64_t row0, uint64_t n,
                          uint32_t dim, float *E, uint64_t *bloom, uint8_t *flags, int threads);

#ifdef __cplusplus
}
#endif
#endif
