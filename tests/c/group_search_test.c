/*
 * group_search_test -- a C host (no Python, no torch) searching one store through several GPU shards.
 *
 *   group_search_test <store> <queries.f32> <nq> <dev,dev,...>
 *
 * Attaches the store once as ONE context (device = first of the list) and once as a group with one shard per listed
 * device (an ordinal may repeat: two shards on one GPU), runs the same searches through both and requires the
 * group's merged answer to be byte-identical to the single context's: single-query scans (top-10 and top-100,
 * with and without a label mask, both orders), a small batch, the exact mode and -- when the store was attached
 * with a bf16 shadow -- a tensor-core batch.  Prints GROUP_OK.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "splinter_gpu.h"

static int fail(const char *what) {
    fprintf(stderr, "group_search_test: %s: %s\n", what, spl_gpu_last_error());
    return 1;
}

int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s <store> <queries.f32> <nq> <dev,dev,...>\n", argv[0]); return 2; }
    const uint32_t nq = (uint32_t)atoi(argv[3]);
    int32_t devs[64];
    uint32_t ndev = 0;
    for (char *tok = strtok(argv[4], ","); tok && ndev < 64; tok = strtok(NULL, ",")) devs[ndev++] = atoi(tok);

    spl_gpu_opts o;
    memset(&o, 0, sizeof o);
    o.struct_size = sizeof o;
    o.device = devs[0];
    o.flags = SPL_OPT_BF16_MIRROR;
    spl_gpu_ctx *one = spl_gpu_attach(argv[1], &o);
    if (!one) return fail("attach");
    spl_gpu_group *grp = spl_gpu_group_attach(argv[1], ndev, devs, &o);
    if (!grp) return fail("group attach");
    spl_gpu_info info;
    spl_gpu_get_info(one, &info);
    uint64_t covered = 0;
    for (uint32_t i = 0; i < spl_gpu_group_size(grp); i++) {
        spl_gpu_info si;
        spl_gpu_get_info(spl_gpu_group_member(grp, i), &si);
        if (si.slot_begin != covered) { fprintf(stderr, "shard %u starts at %llu, expected %llu\n", i, (unsigned long long)si.slot_begin, (unsigned long long)covered); return 1; }
        covered += si.slot_count;
    }
    if (covered != info.total_slots) { fprintf(stderr, "shards cover %llu of %llu slots\n", (unsigned long long)covered, (unsigned long long)info.total_slots); return 1; }

    float *q = (float *)malloc((size_t)nq * info.dim * sizeof(float));
    FILE *fp = fopen(argv[2], "rb");
    if (!fp || fread(q, sizeof(float), (size_t)nq * info.dim, fp) != (size_t)nq * info.dim) { fprintf(stderr, "cannot read %u queries\n", nq); return 1; }
    fclose(fp);

    static const struct { uint32_t nq, k, order, mode; uint64_t mask; } cases[] = {
        { 1, 10, SPL_ORDER_COSINE, SPL_MODE_FAST, 0 },   { 1, 100, SPL_ORDER_COSINE, SPL_MODE_FAST, 0x1 },
        { 1, 32, SPL_ORDER_L2, SPL_MODE_FAST, 0x3 },     { 5, 10, SPL_ORDER_COSINE, SPL_MODE_FAST, 0x2 },
        { 3, 20, SPL_ORDER_COSINE, SPL_MODE_EXACT, 0 },  { 16, 10, SPL_ORDER_COSINE, SPL_MODE_BATCH_BF16, 0 },
        { 16, 10, SPL_ORDER_L2, SPL_MODE_BATCH_BF16, 0x1 } };
    int checked = 0;
    for (size_t ci = 0; ci < sizeof cases / sizeof cases[0]; ci++) {
        if (cases[ci].nq > nq) continue;
        spl_gpu_search_params p;
        memset(&p, 0, sizeof p);
        p.struct_size = sizeof p; p.nq = cases[ci].nq; p.dim = info.dim; p.k = cases[ci].k;
        p.order = cases[ci].order; p.mode = cases[ci].mode; p.bloom_mask = cases[ci].mask; p.exclude_slot = SPL_NO_SLOT;
        const size_t nrec = (size_t)p.nq * p.k;
        spl_gpu_result *a = (spl_gpu_result *)calloc(nrec, sizeof *a), *b = (spl_gpu_result *)calloc(nrec, sizeof *b);
        uint32_t ca[64], cb[64];
        if (spl_gpu_search(one, &p, q, a, p.k, ca) != 0) return fail("single-context search");
        if (spl_gpu_group_search(grp, &p, q, b, p.k, cb) != 0) return fail("group search");
        for (uint32_t i = 0; i < p.nq; i++) {
            if (ca[i] != cb[i]) { fprintf(stderr, "case %zu query %u: %u hits vs %u\n", ci, i, ca[i], cb[i]); return 1; }
            if (ca[i] == 0) { fprintf(stderr, "case %zu query %u: no hits at all\n", ci, i); return 1; }
            if (memcmp(a + (size_t)i * p.k, b + (size_t)i * p.k, (size_t)ca[i] * sizeof *a) != 0) {
                fprintf(stderr, "case %zu query %u: merged records differ from the single-context answer\n", ci, i);
                return 1;
            }
        }
        free(a); free(b);
        checked++;
    }
    /* a write to the store is picked up by every shard's own sweep */
    spl_gpu_sync_stats st;
    if (spl_gpu_group_sync(grp, &st) != 0) return fail("group sync");
    if (st.slots_swept != info.total_slots) { fprintf(stderr, "group sync swept %llu slots\n", (unsigned long long)st.slots_swept); return 1; }
    printf("GROUP_OK shards=%u cases=%d reissued=%llu\n", spl_gpu_group_size(grp), checked, (unsigned long long)spl_gpu_group_reissued(grp));
    spl_gpu_group_detach(grp);
    spl_gpu_detach(one);
    free(q);
    return 0;
}
