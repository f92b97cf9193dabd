/*
 * oracle/ref_backfill.c -- TEST INFRASTRUCTURE ONLY.
 *
 * The reference's way of embedding ingested text, restated around a stub model: splinference's backfill
 * (/root/reference/splinference.cpp:289-325) lists the keys, snapshots each one, skips what is not VARTEXT or already
 * carries a vector, and process_key (:196-287) reads the text through splinter_get_raw_ptr, runs the model on THAT ONE
 * text, writes the vector with splinter_set_embedding and checks that the slot epoch moved by exactly 2.  Everything
 * here goes through the reference's own API; only llama_tokenize / llama_decode are replaced by stub_embed_one.
 *
 *   ref_backfill <store> [label_mask_hex]   ->  prints "<embedded> <failed> <seconds>"
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "splinter.h"
#include "stub_embed.h"

static int needs_embedding(const float *v, size_t n) {   /* splinference.cpp:129-133 */
    double s = 0.0;
    for (size_t i = 0; i < n; i++) s += (double)v[i] * v[i];
    return sqrt(s) < 1e-6;
}

static uint64_t process_key(const char *key) {            /* splinference.cpp:196-287 without the model */
    size_t val_len = 0;
    uint64_t current = splinter_get_epoch(key);
    if (current & 1) return 0;
    uint64_t ptr_epoch = 0;
    const void *raw = splinter_get_raw_ptr(key, &val_len, &ptr_epoch);
    if (!raw || val_len == 0 || ptr_epoch != current) return 0;
    float vec[SPLINTER_EMBED_DIM];
    stub_embed_one((const char *)raw, (uint32_t)val_len, SPLINTER_EMBED_DIM, vec);
    if (splinter_get_epoch(key) != current) return 0;
    if (splinter_set_embedding(key, vec) != 0) return 0;
    uint64_t post = splinter_get_epoch(key);
    return post == current + 2 ? post : 0;
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s <store> [label_mask_hex]\n", argv[0]); return 2; }
    const uint64_t mask = argc > 2 ? strtoull(argv[2], NULL, 16) : 0;
    if (splinter_open(argv[1]) != 0) { perror("splinter_open"); return 1; }
    splinter_header_snapshot_t hs;
    splinter_get_header_snapshot(&hs);
    char **keys = calloc(hs.slots, sizeof(char *));
    size_t n = 0;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    if (splinter_list(keys, hs.slots, &n) != 0) return 1;
    unsigned long embedded = 0, failed = 0;
    for (size_t i = 0; i < n; i++) {
        splinter_slot_snapshot_t snap;
        char key[SPLINTER_KEY_MAX];
        strncpy(key, keys[i], SPLINTER_KEY_MAX - 1);
        key[SPLINTER_KEY_MAX - 1] = '\0';
        if (splinter_get_slot_snapshot(key, &snap) != 0) continue;
        if ((snap.bloom & mask) != mask) continue;
        if ((snap.type_flag & SPL_SLOT_TYPE_VARTEXT) && needs_embedding(snap.embedding, SPLINTER_EMBED_DIM)) {
            if (process_key(key) != 0) embedded++; else failed++;
        }
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    printf("%lu %lu %.6f\n", embedded, failed, (t1.tv_sec - t0.tv_sec) + (t1.tv_nsec - t0.tv_nsec) / 1e9);
    return 0;
}
