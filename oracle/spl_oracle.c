/*
 * oracle/spl_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of libsplinter's similarity-search hot path (see
 * spl_oracle.h for the per-function reference citations).  It is the checker
 * for the CUDA path; it is never shipped, linked or called by the product.
 *
 * Parity status: PINNED against the reference compiled unmodified into
 * oracle/_ref/ (tests/test_oracle_pinned.py) and the committed goldens.
 *
 * Build: oracle/Makefile (`make restatement`): gcc -O3, no -ffast-math, no -march,
 * so fp32 sums keep the reference's sequential rounding.
 */
#include "spl_oracle.h"

#include <errno.h>
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ layout helpers */

typedef struct {
    uint8_t *base;
    uint32_t slots, max_val_sz, dim;
    size_t slot_size;
    uint8_t *S, *V;
} img_t;

static uint32_t rd32(const void *p) { uint32_t v; memcpy(&v, p, 4); return v; }
static uint64_t rd64(const void *p) { uint64_t v; memcpy(&v, p, 8); return v; }
static void wr32(void *p, uint32_t v) { memcpy(p, &v, 4); }
static void wr64(void *p, uint64_t v) { memcpy(p, &v, 8); }

/* header field offsets (splinter.h:176-216, measured) */
enum { H_MAGIC = 0, H_VERSION = 4, H_SLOTS = 8, H_MAXVAL = 12, H_EPOCH = 16, H_CORE = 24, H_USER = 25,
       H_VALBRK = 28, H_VALSZ = 32, H_BLOOMW = 56, H_EVBUS = 4224, H_EVBUS_FD = 4224 + 128 };

static int img_open_dim(img_t *im, const void *base, uint32_t dim) {
    im->base = (uint8_t *)base;
    if (rd32(im->base + H_MAGIC) != SPLO_MAGIC || rd32(im->base + H_VERSION) != SPLO_VERSION) return -1;
    im->slots = rd32(im->base + H_SLOTS);
    im->max_val_sz = rd32(im->base + H_MAXVAL);
    im->dim = dim;
    im->slot_size = SPLO_SLOT_META + 4u * (size_t)dim; /* alignas(64): 128 + 4*dim is a multiple of 64 when dim%16==0 */
    im->S = im->base + SPLO_HEADER_SIZE;
    im->V = im->S + (size_t)im->slots * im->slot_size;
    return 0;
}

/* The image does not record dim; we stash it in the (otherwise unused by create) `alignment`
 * header word?  No: that would make images differ from the reference's.  Instead every
 * image function takes the dim from a process-local registry keyed by base pointer, falling
 * back to 768 (the reference's SPLINTER_EMBED_DIM, splinter.h:39). */
#define REG_MAX 64
static struct { const void *base; uint32_t dim; } g_reg[REG_MAX];
static pthread_mutex_t g_reg_mu = PTHREAD_MUTEX_INITIALIZER;

static void reg_set(const void *base, uint32_t dim) {
    pthread_mutex_lock(&g_reg_mu);
    int free_i = -1;
    for (int i = 0; i < REG_MAX; i++) {
        if (g_reg[i].base == base) { g_reg[i].dim = dim; pthread_mutex_unlock(&g_reg_mu); return; }
        if (!g_reg[i].base && free_i < 0) free_i = i;
    }
    if (free_i < 0) {   /* table full: the oldest entries belong to images freed long ago -- overwrite round-robin */
        static int next_victim = 0;
        free_i = next_victim;
        next_victim = (next_victim + 1) % REG_MAX;
    }
    g_reg[free_i].base = base; g_reg[free_i].dim = dim;
    pthread_mutex_unlock(&g_reg_mu);
}
static uint32_t reg_get(const void *base) {
    uint32_t d = 768;
    pthread_mutex_lock(&g_reg_mu);
    for (int i = 0; i < REG_MAX; i++) if (g_reg[i].base == base) { d = g_reg[i].dim; break; }
    pthread_mutex_unlock(&g_reg_mu);
    return d;
}
static int img_open(img_t *im, const void *base) { return img_open_dim(im, base, reg_get(base)); }

static uint8_t *slot_at(const img_t *im, size_t i) { return im->S + i * im->slot_size; }

/* ------------------------------------------------------------------ scalar pieces */

uint64_t splo_fnv1a(const char *s) { /* splinter.c:67-71 */
    uint64_t h = 14695981039346656037ULL;
    for (; *s; ++s) h = (h ^ (unsigned char)*s) * 1099511628211ULL;
    return h;
}

float splo_cosine(const float *a, const float *b, int dim) { /* _search.c:43-53 */
    float dot = 0.0f, ma = 0.0f, mb = 0.0f;
    for (int i = 0; i < dim; i++) {
        dot += a[i] * b[i];
        ma += a[i] * a[i];
        mb += b[i] * b[i];
    }
    float den = sqrtf(ma) * sqrtf(mb);
    if (den < 1e-9f) return 0.0f;
    return dot / den;
}

float splo_euclid(const float *a, const float *b, int dim) { /* _search.c:55-62 */
    float sum = 0.0f;
    for (int i = 0; i < dim; i++) {
        float d = a[i] - b[i];
        sum += d * d;
    }
    return sqrtf(sum);
}

int splo_compare(int order, float sa, float da, uint32_t ia, float sb, float db, uint32_t ib) {
    if (order == SPLO_ORDER_COSINE) { /* _search.c:64-72 */
        if (sb > sa) return 1;
        if (sb < sa) return -1;
        if (da < db) return -1;
        if (da > db) return 1;
    } else {
        if (da < db) return -1;
        if (da > db) return 1;
        if (sb > sa) return 1;
        if (sb < sa) return -1;
    }
    return ia < ib ? -1 : (ia > ib ? 1 : 0); /* the reference leaves ties unspecified; we state idx ascending */
}

/* ------------------------------------------------------------------ image create / mutate */

size_t splo_image_size(uint32_t slots, uint32_t max_val_sz, uint32_t dim) {
    return SPLO_HEADER_SIZE + (size_t)slots * (SPLO_SLOT_META + 4u * (size_t)dim) + (size_t)slots * max_val_sz;
}

int splo_image_init(void *base, uint32_t slots, uint32_t max_val_sz, uint32_t dim) { /* splinter.c:148-233 */
    if (!slots || !max_val_sz || dim % 16) return -2;
    uint8_t *b = (uint8_t *)base;
    size_t total = splo_image_size(slots, max_val_sz, dim);
    memset(b, 0, total);
    wr32(b + H_MAGIC, SPLO_MAGIC);
    wr32(b + H_VERSION, SPLO_VERSION);
    wr32(b + H_SLOTS, slots);
    wr32(b + H_MAXVAL, max_val_sz);
    wr32(b + H_VALSZ, (uint32_t)total);
    wr64(b + H_EPOCH, 1);
    b[H_CORE] = 0x3; /* SPL_SYS_AUTO_SCRUB | SPL_SYS_HYBRID_SCRUB */
    memset(b + H_BLOOMW, 0xFF, 64);
    wr32(b + H_EVBUS_FD, (uint32_t)-1);
    reg_set(base, dim);
    img_t im;
    img_open_dim(&im, base, dim);
    for (size_t i = 0; i < slots; i++) {
        uint8_t *s = slot_at(&im, i);
        s[SPLO_OFF_TYPE] = 1; /* SPL_SLOT_TYPE_VOID */
        wr32(s + SPLO_OFF_VAL_OFF, (uint32_t)(i * max_val_sz));
    }
    return 0;
}

uint32_t splo_image_dim(const void *base, size_t file_size) {
    const uint8_t *b = (const uint8_t *)base;
    if (file_size < SPLO_HEADER_SIZE || rd32(b + H_MAGIC) != SPLO_MAGIC || rd32(b + H_VERSION) != SPLO_VERSION) return 0;
    uint64_t slots = rd32(b + H_SLOTS), mv = rd32(b + H_MAXVAL);
    if (!slots) return 0;
    uint64_t rest = file_size - SPLO_HEADER_SIZE;
    if (rest < slots * mv) return 0;
    rest -= slots * mv;
    if (rest % slots) return 0;
    uint64_t ss = rest / slots;
    if (ss <= SPLO_SLOT_META || (ss - SPLO_SLOT_META) % 4) return 0;
    uint32_t dim = (uint32_t)((ss - SPLO_SLOT_META) / 4);
    reg_set(base, dim);
    return dim;
}

static long probe(const img_t *im, const char *key, uint64_t h, int for_insert) {
    size_t idx = (size_t)(h % im->slots);
    for (size_t i = 0; i < im->slots; i++) {
        size_t s = (idx + i) % im->slots;
        const uint8_t *sl = slot_at(im, s);
        uint64_t sh = rd64(sl + SPLO_OFF_HASH);
        if ((for_insert && sh == 0) || (sh == h && strncmp((const char *)sl + SPLO_OFF_KEY, key, SPLO_KEY_MAX) == 0))
            return (long)s;
    }
    return -1;
}

long splo_image_find(const void *base, const char *key) {
    img_t im;
    if (img_open(&im, base)) return -1;
    return probe(&im, key, splo_fnv1a(key), 0);
}

static void write_slot_value(img_t *im, uint8_t *sl, int fresh, const char *key, uint64_t h, const void *val, uint32_t len) {
    uint8_t *dst = im->V + rd32(sl + SPLO_OFF_VAL_OFF);
    size_t scrub = ((size_t)len + 63) & ~(size_t)63; /* hybrid scrub, splinter.c:392-396 */
    if (scrub > im->max_val_sz) scrub = im->max_val_sz;
    memset(dst, 0, scrub);
    memcpy(dst, val, len);
    wr32(sl + SPLO_OFF_VAL_LEN, len);
    if (fresh) memset(sl + SPLO_OFF_EMBED, 0, 4u * (size_t)im->dim); /* splinter.c:406-411 */
    char *k = (char *)sl + SPLO_OFF_KEY;
    k[0] = '\0';
    strncpy(k, key, SPLO_KEY_MAX - 1);
    k[SPLO_KEY_MAX - 1] = '\0';
    wr64(sl + SPLO_OFF_HASH, h);
    wr64(sl + SPLO_OFF_EPOCH, rd64(sl + SPLO_OFF_EPOCH) + 2);
}

long splo_image_set(void *base, const char *key, const void *val, uint32_t len) {
    img_t im;
    if (img_open(&im, base)) return -2;
    if (len == 0 || len > im.max_val_sz) return -1;
    uint64_t h = splo_fnv1a(key);
    long s = probe(&im, key, h, 1);
    if (s < 0) return -1;
    uint8_t *sl = slot_at(&im, (size_t)s);
    write_slot_value(&im, sl, rd64(sl + SPLO_OFF_HASH) == 0, key, h, val, len);
    wr64(im.base + H_EPOCH, rd64(im.base + H_EPOCH) + 1);
    return s;
}

int splo_image_set_embedding(void *base, const char *key, const float *vec) {
    img_t im;
    if (img_open(&im, base)) return -2;
    long s = probe(&im, key, splo_fnv1a(key), 0);
    if (s < 0) return -1;
    uint8_t *sl = slot_at(&im, (size_t)s);
    memcpy(sl + SPLO_OFF_EMBED, vec, 4u * (size_t)im.dim);
    wr64(sl + SPLO_OFF_EPOCH, rd64(sl + SPLO_OFF_EPOCH) + 2);
    wr64(im.base + H_EPOCH, rd64(im.base + H_EPOCH) + 1);
    return 0;
}

static int label_op(void *base, const char *key, uint64_t mask, int set) {
    img_t im;
    if (img_open(&im, base)) return -2;
    long s = probe(&im, key, splo_fnv1a(key), 0);
    if (s < 0) return -1;
    uint8_t *sl = slot_at(&im, (size_t)s);
    uint64_t b = rd64(sl + SPLO_OFF_BLOOM);
    wr64(sl + SPLO_OFF_BLOOM, set ? (b | mask) : (b & ~mask));
    wr64(im.base + H_EPOCH, rd64(im.base + H_EPOCH) + 1);
    return 0;
}
int splo_image_set_label(void *base, const char *key, uint64_t mask) { return label_op(base, key, mask, 1); }
int splo_image_unset_label(void *base, const char *key, uint64_t mask) { return label_op(base, key, mask, 0); }

int splo_image_retrain(void *base, const char *key) {
    img_t im;
    if (img_open(&im, base)) return -2;
    long s = probe(&im, key, splo_fnv1a(key), 0);
    if (s < 0) return -1;
    uint8_t *sl = slot_at(&im, (size_t)s);
    memset(sl + SPLO_OFF_EMBED, 0, 4u * (size_t)im.dim);
    wr64(sl + SPLO_OFF_EPOCH, 4); /* may move backwards, splinter.c:807-825 */
    wr64(im.base + H_EPOCH, rd64(im.base + H_EPOCH) + 1);
    return 0;
}

int splo_image_unset(void *base, const char *key) {
    img_t im;
    if (img_open(&im, base)) return -2;
    long s = probe(&im, key, splo_fnv1a(key), 0);
    if (s < 0) return -1;
    uint8_t *sl = slot_at(&im, (size_t)s);
    int ret = (int)rd32(sl + SPLO_OFF_VAL_LEN);
    uint32_t val_off = rd32(sl + SPLO_OFF_VAL_OFF);
    memset(im.V + val_off, 0, im.max_val_sz);
    memset(sl, 0, im.slot_size); /* hash, key, lens, times, flags, bloom, embedding -> 0 */
    wr32(sl + SPLO_OFF_VAL_OFF, val_off);
    sl[SPLO_OFF_TYPE] = 1;
    wr64(sl + SPLO_OFF_EPOCH, 2); /* := 0 then += 2, splinter.c:350,358; global epoch untouched */
    return ret;
}

/* ------------------------------------------------------------------ synthetic data */

uint64_t splo_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

int splo_synth_class(uint64_t seed, uint64_t row) {
    uint64_t c = splo_mix64(splo_mix64(seed ^ 0xC1A55C1A55ULL) ^ (row * 0xD1342543DE82EF95ULL)) % 1000u;
    if (c < 1) return SPLO_CLASS_EMPTY;
    if (c < 21) return SPLO_CLASS_ZERO;
    return SPLO_CLASS_EMBED;
}

void splo_synth_row(uint64_t seed, uint64_t row, uint32_t dim, float *out) {
    uint64_t hr = splo_mix64(seed ^ (row * 0x9E3779B97F4A7C15ULL));
    float ss = 0.0f;
    for (uint32_t d = 0; d < dim; d++) {
        uint64_t h = splo_mix64(hr + d);
        int32_t m = (int32_t)(h >> 40) - (1 << 23);  /* 24 random bits -> [-2^23, 2^23) */
        float v = (float)m * (1.0f / 8388608.0f);    /* exact: uniform grid on [-1, 1) */
        out[d] = v;
        ss += v * v;                                  /* sequential fp32, like the GPU generator */
    }
    float inv = 1.0f / sqrtf(ss);
    for (uint32_t d = 0; d < dim; d++) out[d] = out[d] * inv;
}

uint64_t splo_synth_bloom(uint64_t seed, uint64_t row) {
    uint64_t hb = splo_mix64(seed ^ (row * 0xA24BAED4963EE407ULL));
    uint64_t bloom = 1ULL << 63;
    for (unsigned b = 0; b < 63; b++) {
        uint64_t h = splo_mix64(hb + b);
        unsigned nb = 1 + (b & 3);
        if ((h & ((1u << nb) - 1)) == 0) bloom |= 1ULL << b;
    }
    return bloom;
}

typedef struct {
    uint64_t seed_c, seed_b, row0, lo, hi;
    uint32_t dim;
    float *E; uint64_t *bloom; uint8_t *flags;
} synth_job_t;

static void *synth_run(void *p) {
    synth_job_t *j = (synth_job_t *)p;
    for (uint64_t r = j->lo; r < j->hi; r++) {
        uint64_t g = j->row0 + r;
        int c = splo_synth_class(j->seed_c, g);
        float *row = j->E + r * (uint64_t)j->dim;
        if (c == SPLO_CLASS_EMBED) splo_synth_row(j->seed_c, g, j->dim, row);
        else memset(row, 0, 4u * (size_t)j->dim);
        j->bloom[r] = c == SPLO_CLASS_EMPTY ? 0 : splo_synth_bloom(j->seed_b, g);
        j->flags[r] = c == SPLO_CLASS_EMPTY ? 0 : (SPLO_F_LIST | SPLO_F_ENUM | SPLO_F_EMB);
    }
    return NULL;
}

static void run_jobs(void *(*fn)(void *), void *jobs, size_t jsz, int threads) {
    if (threads <= 1) { fn(jobs); return; }
    pthread_t *t = (pthread_t *)calloc((size_t)threads, sizeof *t);
    for (int i = 0; i < threads; i++) pthread_create(&t[i], NULL, fn, (uint8_t *)jobs + (size_t)i * jsz);
    for (int i = 0; i < threads; i++) pthread_join(t[i], NULL);
    free(t);
}

void splo_synth_dense(uint64_t seed_c, uint64_t seed_b, uint64_t row0, uint64_t n, uint32_t dim,
                      float *E, uint64_t *bloom, uint8_t *flags, int threads) {
    if (threads < 1) threads = 1;
    synth_job_t *jobs = (synth_job_t *)calloc((size_t)threads, sizeof *jobs);
    uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    for (int t = 0; t < threads; t++) {
        uint64_t lo = per * (uint64_t)t, hi = lo + per;
        if (lo > n) lo = n;
        if (hi > n) hi = n;
        jobs[t] = (synth_job_t){ seed_c, seed_b, row0, lo, hi, dim, E, bloom, flags };
    }
    run_jobs(synth_run, jobs, sizeof *jobs, threads);
    free(jobs);
}

typedef struct {
    img_t im;
    const uint32_t *slot_of_key;
    uint64_t seed_c, seed_b;
    uint32_t lo, hi;
    uint64_t n_emb;
} fill_job_t;

static void *fill_run(void *p) {
    fill_job_t *j = (fill_job_t *)p;
    for (uint32_t k = j->lo; k < j->hi; k++) {
        uint32_t s = j->slot_of_key[k];
        if (s == UINT32_MAX) continue;
        char key[SPLO_KEY_MAX], val[16];
        snprintf(key, sizeof key, "doc%08u", k);
        int vl = snprintf(val, sizeof val, "v%08u", k);
        uint8_t *sl = slot_at(&j->im, s);
        write_slot_value(&j->im, sl, 1, key, splo_fnv1a(key), val, (uint32_t)vl); /* epoch 0 -> 2 */
        if (splo_synth_class(j->seed_c, k) == SPLO_CLASS_EMBED) {
            splo_synth_row(j->seed_c, k, j->im.dim, (float *)(sl + SPLO_OFF_EMBED));
            wr64(sl + SPLO_OFF_EPOCH, 4);                                          /* set_embedding: 2 -> 4 */
            j->n_emb++;
        }
        wr64(sl + SPLO_OFF_BLOOM, splo_synth_bloom(j->seed_b, k));
    }
    return NULL;
}

long splo_image_fill_synthetic(void *base, uint32_t nkeys, uint64_t seed_c, uint64_t seed_b, int threads) {
    img_t im;
    if (img_open(&im, base)) return -2;
    if (threads < 1) threads = 1;
    /* phase 1 (serial): linear-probe placement exactly as splinter_set would do it, using a
     * compact occupancy map so a 99.9 %-full table stays fast */
    uint8_t *occ = (uint8_t *)calloc(im.slots, 1);
    uint32_t *slot_of_key = (uint32_t *)malloc((size_t)nkeys * sizeof(uint32_t));
    if (!occ || !slot_of_key) { free(occ); free(slot_of_key); return -1; }
    for (size_t i = 0; i < im.slots; i++) occ[i] = rd64(slot_at(&im, i) + SPLO_OFF_HASH) != 0;
    long placed = 0;
    for (uint32_t k = 0; k < nkeys; k++) {
        slot_of_key[k] = UINT32_MAX;
        if (splo_synth_class(seed_c, k) == SPLO_CLASS_EMPTY) continue;
        char key[SPLO_KEY_MAX];
        snprintf(key, sizeof key, "doc%08u", k);
        size_t idx = (size_t)(splo_fnv1a(key) % im.slots);
        for (size_t i = 0; i < im.slots; i++) {
            size_t s = idx + i;
            if (s >= im.slots) s -= im.slots;
            if (!occ[s]) { occ[s] = 1; slot_of_key[k] = (uint32_t)s; placed++; break; }
        }
    }
    fill_job_t *jobs = (fill_job_t *)calloc((size_t)threads, sizeof *jobs);
    uint32_t per = (nkeys + (uint32_t)threads - 1) / (uint32_t)threads;
    for (int t = 0; t < threads; t++) {
        uint64_t lo = (uint64_t)per * (uint64_t)t, hi = lo + per;
        if (lo > nkeys) lo = nkeys;
        if (hi > nkeys) hi = nkeys;
        jobs[t] = (fill_job_t){ im, slot_of_key, seed_c, seed_b, (uint32_t)lo, (uint32_t)hi, 0 };
    }
    run_jobs(fill_run, jobs, sizeof *jobs, threads);
    uint64_t n_emb = 0;
    for (int t = 0; t < threads; t++) n_emb += jobs[t].n_emb;
    /* global epoch: +1 per set, per set_embedding, per set_label (one call per key) */
    wr64(im.base + H_EPOCH, rd64(im.base + H_EPOCH) + 2u * (uint64_t)placed + n_emb);
    free(jobs); free(occ); free(slot_of_key);
    return placed;
}

/* ------------------------------------------------------------------ candidates / snapshot */

size_t splo_enumerate_matches(const void *base, uint64_t mask, uint32_t *out_idx) { /* splinter.c:977-989 */
    img_t im;
    if (img_open(&im, base)) return 0;
    size_t n = 0;
    for (size_t i = 0; i < im.slots; i++) {
        const uint8_t *sl = slot_at(&im, i);
        if (rd64(sl + SPLO_OFF_HASH) != 0 && (rd64(sl + SPLO_OFF_BLOOM) & mask) == mask) out_idx[n++] = (uint32_t)i;
    }
    return n;
}

size_t splo_list(const void *base, uint32_t *out_idx) { /* splinter.c:466-477 */
    img_t im;
    if (img_open(&im, base)) return 0;
    size_t n = 0;
    for (size_t i = 0; i < im.slots; i++) {
        const uint8_t *sl = slot_at(&im, i);
        if (rd64(sl + SPLO_OFF_HASH) != 0 && rd32(sl + SPLO_OFF_VAL_LEN) > 0) out_idx[n++] = (uint32_t)i;
    }
    return n;
}

int splo_snapshot_row(const void *base, uint32_t idx, float *row_out, uint64_t *epoch_out) { /* splinter.c:598-605 */
    img_t im;
    if (img_open(&im, base)) return -2;
    const uint8_t *sl = slot_at(&im, idx);
    uint64_t e1 = __atomic_load_n((const uint64_t *)(sl + SPLO_OFF_EPOCH), __ATOMIC_ACQUIRE);
    if (e1 & 1) { errno = EAGAIN; return -1; }
    memcpy(row_out, sl + SPLO_OFF_EMBED, 4u * (size_t)im.dim);
    __atomic_thread_fence(__ATOMIC_ACQUIRE);
    uint64_t e2 = __atomic_load_n((const uint64_t *)(sl + SPLO_OFF_EPOCH), __ATOMIC_ACQUIRE);
    if (epoch_out) *epoch_out = e1;
    if (e1 != e2) { errno = EAGAIN; return -1; }
    return 0;
}

static uint8_t slot_flags(const uint8_t *sl) {
    uint64_t h = rd64(sl + SPLO_OFF_HASH), e = rd64(sl + SPLO_OFF_EPOCH);
    uint32_t vl = rd32(sl + SPLO_OFF_VAL_LEN);
    int key0 = sl[SPLO_OFF_KEY] != 0;
    uint8_t f = 0;
    if (h && e && key0) f |= SPLO_F_ENUM;
    if (h && e && key0 && vl > 0) f |= SPLO_F_LIST;
    if (!(e & 1)) f |= SPLO_F_EMB;
    return f;
}

void splo_image_to_dense(const void *base, float *E, uint64_t *bloom, uint8_t *flags, uint64_t *epoch) {
    img_t im;
    if (img_open(&im, base)) return;
    for (size_t i = 0; i < im.slots; i++) {
        const uint8_t *sl = slot_at(&im, i);
        if (E) memcpy(E + i * (size_t)im.dim, sl + SPLO_OFF_EMBED, 4u * (size_t)im.dim);
        if (bloom) bloom[i] = rd64(sl + SPLO_OFF_BLOOM);
        if (flags) flags[i] = slot_flags(sl);
        if (epoch) epoch[i] = rd64(sl + SPLO_OFF_EPOCH);
    }
}

/* ------------------------------------------------------------------ search */

static int g_order; /* qsort context (searches are serialised per thread-pool join; see below) */
static int res_cmp(const void *a, const void *b) {
    const splo_result *x = (const splo_result *)a, *y = (const splo_result *)b;
    return splo_compare(g_order, x->similarity, x->distance, x->idx, y->similarity, y->distance, y->idx);
}
static int res_cmp_order(const splo_result *x, const splo_result *y, int order) {
    return splo_compare(order, x->similarity, x->distance, x->idx, y->similarity, y->distance, y->idx);
}

/* bounded insertion into a sorted top-`limit` array (stable total order, so the outcome equals
 * a full qsort followed by truncation) */
static void topk_push(splo_result *top, uint32_t *cnt, uint32_t limit, const splo_result *r, int order) {
    if (*cnt == limit && res_cmp_order(r, &top[limit - 1], order) >= 0) return;
    uint32_t lo = 0, hi = *cnt;
    while (lo < hi) {
        uint32_t mid = (lo + hi) / 2;
        if (res_cmp_order(&top[mid], r, order) <= 0) lo = mid + 1; else hi = mid;
    }
    uint32_t last = *cnt < limit ? *cnt : limit - 1;
    memmove(&top[lo + 1], &top[lo], (size_t)(last - lo) * sizeof *top);
    top[lo] = *r;
    if (*cnt < limit) (*cnt)++;
}

static int score_candidate(const float *q, const float *row, uint32_t dim, int has_emb,
                           float min_sim, float max_dist, splo_result *r) { /* _search.c:385-400 */
    if (!has_emb) {
        r->similarity = 0.0f; r->distance = FLT_MAX; r->has_embedding = 0;
        if (min_sim > 0.0f || max_dist > 0.0f) return 0;
        return 1;
    }
    r->similarity = splo_cosine(q, row, (int)dim);
    r->distance = splo_euclid(q, row, (int)dim);
    r->has_embedding = 1;
    if (min_sim > 0.0f && r->similarity < min_sim) return 0;
    if (max_dist > 0.0f && r->distance > max_dist) return 0;
    return 1;
}

long splo_search_image(const void *base, const float *q, int order, int limit, uint64_t bloom_mask,
                       float min_sim, float max_dist, const char *exclude_key,
                       splo_result *out, long *survivors) {
    img_t im;
    if (img_open(&im, base)) return -2;
    uint32_t *cand = (uint32_t *)malloc((size_t)im.slots * sizeof *cand);
    splo_result *res = (splo_result *)malloc((size_t)im.slots * sizeof *res);
    float *row = (float *)malloc(4u * (size_t)im.dim);
    if (!cand || !res || !row) { free(cand); free(res); free(row); return -1; }
    size_t nc = bloom_mask ? splo_enumerate_matches(base, bloom_mask, cand) : splo_list(base, cand); /* :337-347 */
    long n = 0;
    for (size_t c = 0; c < nc; c++) {
        const uint8_t *sl = slot_at(&im, cand[c]);
        const char *key = (const char *)sl + SPLO_OFF_KEY;
        if (key[0] == '\0') continue;                                              /* :366 */
        if (exclude_key && strncmp(key, exclude_key, SPLO_KEY_MAX) == 0) continue; /* :367 */
        if (rd64(sl + SPLO_OFF_EPOCH) == 0) continue;                              /* :375 */
        int has = splo_snapshot_row(base, cand[c], row, NULL) == 0;                /* :387 */
        splo_result r; r.idx = cand[c];
        if (score_candidate(q, row, im.dim, has, min_sim, max_dist, &r)) res[n++] = r;
    }
    g_order = order;
    qsort(res, (size_t)n, sizeof *res, res_cmp);                                   /* :412 (+ idx tie-break) */
    long outn = n;
    if (limit > 0 && outn > limit) outn = limit;                                   /* :415-416 */
    memcpy(out, res, (size_t)outn * sizeof *res);
    if (survivors) *survivors = n;
    free(cand); free(res); free(row);
    return outn;
}

typedef struct {
    const float *q, *E; const uint64_t *bloom; const uint8_t *flags;
    uint64_t lo, hi; uint32_t dim; int order; uint32_t limit; uint64_t mask;
    float min_sim, max_dist; uint32_t exclude;
    splo_result *top; uint32_t cnt; uint64_t survivors;
} dense_job_t;

static void *dense_run(void *p) {
    dense_job_t *j = (dense_job_t *)p;
    uint8_t need = j->mask ? SPLO_F_ENUM : SPLO_F_LIST;
    j->cnt = 0; j->survivors = 0;
    for (uint64_t i = j->lo; i < j->hi; i++) {
        uint8_t f = j->flags ? j->flags[i] : (uint8_t)(SPLO_F_LIST | SPLO_F_ENUM | SPLO_F_EMB);
        if (!(f & need)) continue;
        if (j->mask && ((j->bloom ? j->bloom[i] : 0) & j->mask) != j->mask) continue;
        if ((uint32_t)i == j->exclude) continue;
        splo_result r; r.idx = (uint32_t)i;
        if (!score_candidate(j->q, j->E + i * (uint64_t)j->dim, j->dim, (f & SPLO_F_EMB) != 0, j->min_sim, j->max_dist, &r)) continue;
        j->survivors++;
        topk_push(j->top, &j->cnt, j->limit, &r, j->order);
    }
    return NULL;
}

long splo_search_dense(const float *q, uint32_t nq, const float *E, const uint64_t *bloom,
                       const uint8_t *flags, uint64_t n, uint32_t dim, int order, int limit,
                       uint64_t bloom_mask, float min_sim, float max_dist, uint32_t exclude_idx,
                       int threads, splo_result *out, uint32_t *counts) {
    if (threads < 1) threads = 1;
    uint32_t lim = limit > 0 ? (uint32_t)limit : (uint32_t)n;
    if (lim == 0) lim = 1;
    dense_job_t *jobs = (dense_job_t *)calloc((size_t)threads, sizeof *jobs);
    splo_result *tops = (splo_result *)malloc((size_t)threads * lim * sizeof *tops);
    if (!jobs || !tops) { free(jobs); free(tops); return -1; }
    uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    long total_survivors = 0;
    for (uint32_t qi = 0; qi < nq; qi++) {
        for (int t = 0; t < threads; t++) {
            uint64_t lo = per * (uint64_t)t, hi = lo + per;
            if (lo > n) lo = n;
            if (hi > n) hi = n;
            jobs[t] = (dense_job_t){ q + (size_t)qi * dim, E, bloom, flags, lo, hi, dim, order, lim, bloom_mask,
                                     min_sim, max_dist, exclude_idx, tops + (size_t)t * lim, 0, 0 };
        }
        run_jobs(dense_run, jobs, sizeof *jobs, threads);
        splo_result *o = out + (size_t)qi * lim;
        uint32_t cnt = 0;
        for (int t = 0; t < threads; t++) {
            total_survivors += (long)jobs[t].survivors;
            for (uint32_t i = 0; i < jobs[t].cnt; i++) topk_push(o, &cnt, lim, &jobs[t].top[i], order);
        }
        if (counts) counts[qi] = cnt;
    }
    free(jobs); free(tops);
    return total_survivors;
}

uint32_t splo_merge_topk(const splo_result *lists, const uint32_t *counts, uint32_t nlists,
                         uint32_t stride, int order, uint32_t k, splo_result *out) {
    uint32_t cnt = 0;
    for (uint32_t l = 0; l < nlists; l++)
        for (uint32_t i = 0; i < counts[l]; i++) topk_push(out, &cnt, k, &lists[(size_t)l * stride + i], order);
    return cnt;
}
