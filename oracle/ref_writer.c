/*
 * oracle/ref_writer.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A writer process that mutates a live store through the reference's own API
 * (splinter_set_embedding / splinter_retrain_slot / splinter_set_label /
 * splinter_unset_label / splinter_unset + splinter_set; splinter.c:567-588,
 * 799-833, 835-867, 326-429), in the style of the reference's MRMW stress tool
 * (splinter_chi_sao.c:400-418: each writer owns a disjoint key lane).
 *
 * Every vector it writes is self-describing so a torn mirror row can be
 * detected afterwards: v[0] = key number, v[1] = version, and
 * v[d] = version + d/1024 for d >= 2 (all derived from ONE version).
 *
 *   ref_writer <store> <key_lo> <key_hi> <seed> <duration_ms> [mix: e r l u in %]
 * Keys are "doc%08d".  Prints "<ops> <set_embedding> <retrain> <label> <unset_set> <eagain>".
 *
 * With REF_WRITER_UNIT=1 in the environment every vector is instead a UNIT vector
 * cos(phi) u + sin(phi) w (u, w fixed orthonormal, phi from the version), so that for a unit
 * query any search hit must satisfy distance^2 = 2 - 2 similarity (or be the zero row a
 * retrain / fresh set leaves: similarity 0, distance 1).  A row stitched together from two
 * versions has a different norm and breaks that identity: the check works on SCORES alone,
 * which is all a search returns.
 */
#include <errno.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "splinter.h"

static uint64_t rng_state;
static uint64_t rng(void) { /* splitmix64 */
    uint64_t z = (rng_state += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6;
}

int main(int argc, char **argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s <store> <key_lo> <key_hi> <seed> <duration_ms> [e r l u]\n", argv[0]);
        return 2;
    }
    long lo = atol(argv[2]), hi = atol(argv[3]);
    rng_state = strtoull(argv[4], NULL, 10);
    double dur = atof(argv[5]);
    int pe = argc > 6 ? atoi(argv[6]) : 80, pr = argc > 7 ? atoi(argv[7]) : 10;
    int pl = argc > 8 ? atoi(argv[8]) : 5; /* remainder: unset+set */
    if (splinter_open(argv[1]) != 0) { perror("splinter_open"); return 1; }

    float vec[SPLINTER_EMBED_DIM];
    const char *unit_env = getenv("REF_WRITER_UNIT");
    const int unit = unit_env && unit_env[0] == '1';
    static double U[SPLINTER_EMBED_DIM], W[SPLINTER_EMBED_DIM];
    if (unit) {   /* fixed orthonormal pair, the same in every writer process */
        uint64_t keep = rng_state;
        rng_state = 777;
        double uu = 0, uw = 0, ww = 0;
        for (int d = 0; d < SPLINTER_EMBED_DIM; d++) {
            U[d] = (double)(rng() >> 11) / 9007199254740992.0 - 0.5;
            W[d] = (double)(rng() >> 11) / 9007199254740992.0 - 0.5;
            uu += U[d] * U[d];
        }
        for (int d = 0; d < SPLINTER_EMBED_DIM; d++) { U[d] /= sqrt(uu); uw += U[d] * W[d]; }
        for (int d = 0; d < SPLINTER_EMBED_DIM; d++) { W[d] -= uw * U[d]; ww += W[d] * W[d]; }
        for (int d = 0; d < SPLINTER_EMBED_DIM; d++) W[d] /= sqrt(ww);
        rng_state = keep;
    }
    unsigned long ops = 0, ne = 0, nr = 0, nl = 0, nu = 0, eagain = 0;
    uint32_t version = 1;
    double t_end = now_ms() + dur;
    while (now_ms() < t_end) {
        for (int burst = 0; burst < 64; burst++) {
            long k = lo + (long)(rng() % (uint64_t)(hi - lo));
            char key[SPLINTER_KEY_MAX];
            snprintf(key, sizeof key, "doc%08ld", k);
            int dice = (int)(rng() % 100);
            int rc = 0;
            if (dice < pe) {
                version++;
                if (unit) {
                    const double phi = 0.61803398875 * (double)version + 0.001 * (double)k;
                    const double c = cos(phi), sn = sin(phi);
                    for (int d = 0; d < SPLINTER_EMBED_DIM; d++) vec[d] = (float)(c * U[d] + sn * W[d]);
                } else {
                    vec[0] = (float)k; vec[1] = (float)version;
                    for (int d = 2; d < SPLINTER_EMBED_DIM; d++) vec[d] = (float)version + (float)d / 1024.0f;
                }
                rc = splinter_set_embedding(key, vec); ne++;
            } else if (dice < pe + pr) {
                rc = splinter_retrain_slot(key); nr++;
            } else if (dice < pe + pr + pl) {
                uint64_t bit = 1ULL << (8 + rng() % 8);
                rc = (rng() & 1) ? splinter_set_label(key, bit) : splinter_unset_label(key, bit); nl++;
            } else {
                splinter_unset(key);
                rc = splinter_set(key, "x", 1); nu++;
            }
            if (rc != 0) eagain++;
            ops++;
        }
    }
    printf("%lu %lu %lu %lu %lu %lu\n", ops, ne, nr, nl, nu, eagain);
    return 0;
}
