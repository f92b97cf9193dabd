/*
 * oracle/ref_harness.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A translation unit that #includes the reference's own search command
 * (/root/reference/splinter_cli_cmd_search.c, found through -I at build time;
 * nothing is copied into this repo) so that its file-static scorers and
 * comparator -- cosine_similarity (:43-53), euclidean_distance (:55-62),
 * compare_results (:64-72) -- and cmd_search (:136-486) itself can be called
 * at full float precision from tests and from bench.py's CPU baseline.
 *
 * Built by oracle/Makefile into oracle/_ref/libref_search.so.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.
 */
#include <pthread.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "splinter_cli_cmd_search.c" /* the reference TU, unmodified */

/* Symbols the reference TU expects from the rest of its CLI. */
cli_user_t thisuser;

char *cli_show_key_type(unsigned short flags) {
    /* label table of splinter_cli_util.c:128-139 (names are the reference's public macros) */
    static const struct { unsigned short bit; const char *name; } t[] = {
        { SPL_SLOT_TYPE_BIGINT, "SPL_SLOT_TYPE_BIGINT" },   { SPL_SLOT_TYPE_BIGUINT, "SPL_SLOT_TYPE_BIGUINT" },
        { SPL_SLOT_TYPE_BINARY, "SPL_SLOT_TYPE_BINARY" },   { SPL_SLOT_TYPE_IMGDATA, "SPL_SLOT_TYPE_IMGDATA" },
        { SPL_SLOT_TYPE_VARTEXT, "SPL_SLOT_TYPE_VARTEXT" }, { SPL_SLOT_TYPE_AUDIO, "SPL_SLOT_TYPE_AUDIO" },
        { SPL_SLOT_TYPE_JSON, "SPL_SLOT_TYPE_JSON" },       { SPL_SLOT_TYPE_VOID, "SPL_SLOT_TYPE_VOID" },
    };
    for (size_t i = 0; i < sizeof t / sizeof t[0]; i++)
        if (flags & t[i].bit) return (char *)t[i].name;
    return "UNNAMED";
}

/* ---- direct access to the static functions -------------------------------- */

float ref_cosine_similarity(const float *a, const float *b, int dim) { return cosine_similarity(a, b, dim); }
float ref_euclidean_distance(const float *a, const float *b, int dim) { return euclidean_distance(a, b, dim); }

int ref_compare(float sim_a, float dist_a, float sim_b, float dist_b) {
    search_result_t a, b;
    memset(&a, 0, sizeof a);
    memset(&b, 0, sizeof b);
    a.similarity = sim_a; a.distance = dist_a;
    b.similarity = sim_b; b.distance = dist_b;
    return compare_results(&a, &b);
}

int ref_embed_dim(void) { return SPLINTER_EMBED_DIM; }
size_t ref_sizeof_slot(void) { return sizeof(struct splinter_slot); }
size_t ref_sizeof_header(void) { return sizeof(struct splinter_header); }
size_t ref_sizeof_result(void) { return sizeof(search_result_t); }
/* where a watcher finds the owner's eventfd (splinter.h:144-148, :210) and the epoch gate (:186) */
size_t ref_offsetof_bus_owner_fd(void) { return offsetof(struct splinter_header, event_bus) + offsetof(struct splinter_event_bus, owner_fd); }
size_t ref_offsetof_bus_owner_pid(void) { return offsetof(struct splinter_header, event_bus) + offsetof(struct splinter_event_bus, owner_pid); }
size_t ref_offsetof_header_epoch(void) { return offsetof(struct splinter_header, epoch); }

/* Run the real command in-process (caller must have splinter_open()ed the
 * store through the same libsplinter.so).  Output goes to stdout as usual. */
int ref_cmd_search(int argc, char **argv) {
    thisuser.store_conn = true;
    return cmd_search(argc, argv);
}

/* ---- scorer + qsort over a dense copy ---------------
 * For each candidate row (cand==NULL: all rows) call the reference scorers,
 * apply the reference's threshold rules (:398-399), qsort with the reference
 * comparator (:412) and keep the first `limit` (:415-416).  The slot index is
 * carried in the record's key field as a decimal string so the unmodified
 * 104-byte search_result_t is what gets sorted. */

typedef struct {
    const float *q, *E;
    const uint8_t *cand;
    size_t lo, hi;
    int dim;
    float min_sim, max_dist;
    search_result_t *res;
    size_t n;
} ref_job_t;

static void *ref_job_run(void *p) {
    ref_job_t *j = (ref_job_t *)p;
    size_t n = 0;
    for (size_t i = j->lo; i < j->hi; i++) {
        if (j->cand && !j->cand[i]) continue;
        const float *row = j->E + i * (size_t)j->dim;
        search_result_t r;
        memset(&r, 0, sizeof r);
        r.similarity = cosine_similarity(j->q, row, j->dim);
        r.distance = euclidean_distance(j->q, row, j->dim);
        r.has_embedding = 1;
        if (j->min_sim > 0.0f && r.similarity < j->min_sim) continue;
        if (j->max_dist > 0.0f && r.distance > j->max_dist) continue;
        snprintf(r.key, sizeof r.key, "%zu", i);
        j->res[n++] = r;
    }
    j->n = n;
    return NULL;
}

/* returns number of survivors (before limit); writes min(limit,survivors) rows */
long ref_topk_dense(const float *q, const float *E, size_t n, int dim, const uint8_t *cand,
                    float min_sim, float max_dist, int limit, int threads,
                    uint32_t *out_idx, float *out_sim, float *out_dist) {
    if (threads < 1) threads = 1;
    search_result_t *all = (search_result_t *)malloc((n ? n : 1) * sizeof *all);
    if (!all) return -1;
    ref_job_t *jobs = (ref_job_t *)calloc((size_t)threads, sizeof *jobs);
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof *tid);
    size_t per = (n + (size_t)threads - 1) / (size_t)threads;
    for (int t = 0; t < threads; t++) {
        size_t lo = per * (size_t)t, hi = lo + per;
        if (lo > n) lo = n;
        if (hi > n) hi = n;
        jobs[t] = (ref_job_t){ q, E, cand, lo, hi, dim, min_sim, max_dist, all + lo, 0 };
        if (threads > 1) pthread_create(&tid[t], NULL, ref_job_run, &jobs[t]);
        else ref_job_run(&jobs[t]);
    }
    size_t total = 0;
    for (int t = 0; t < threads; t++) {
        if (threads > 1) pthread_join(tid[t], NULL);
        /* per-thread partial sort then keep its first `limit`: any global top-`limit`
         * row is inside its own thread's top-`limit` */
        qsort(jobs[t].res, jobs[t].n, sizeof(search_result_t), compare_results);
        size_t keep = jobs[t].n;
        if (threads > 1 && limit > 0 && keep > (size_t)limit) keep = (size_t)limit;
        memmove(all + total, jobs[t].res, keep * sizeof(search_result_t));
        total += keep;
    }
    long survivors = 0;
    for (int t = 0; t < threads; t++) survivors += (long)jobs[t].n;
    qsort(all, total, sizeof(search_result_t), compare_results);
    size_t outn = total;
    if (limit > 0 && outn > (size_t)limit) outn = (size_t)limit;
    for (size_t i = 0; i < outn; i++) {
        out_idx[i] = (uint32_t)strtoul(all[i].key, NULL, 10);
        out_sim[i] = all[i].similarity;
        out_dist[i] = all[i].distance;
    }
    free(tid); free(jobs); free(all);
    return survivors;
}

/* plain scoring of every row (no sort) */
void ref_score_dense(const float *q, const float *E, size_t n, int dim, float *sim, float *dist) {
    for (size_t i = 0; i < n; i++) {
        sim[i] = cosine_similarity(q, E + i * (size_t)dim, dim);
        dist[i] = euclidean_distance(q, E + i * (size_t)dim, dim);
    }
}
