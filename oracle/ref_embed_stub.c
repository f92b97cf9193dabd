/*
 * oracle/ref_embed_stub.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Stands in for the `splinference` daemon (llama.cpp; not buildable here) for ANY search client: it
 * watches the store for a scratch key `__sqtmp_<pid>` carrying label 0x1 (the handshake of
 * /root/reference/splinter_cli_cmd_search.c:301-311) and answers it with splinter_set_embedding(vector
 * read from a file), through the reference's own API.
 *
 *   ref_embed_stub <store> <query.f32> [timeout_ms]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "splinter.h"

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s <store> <query.f32> [timeout_ms]\n", argv[0]); return 2; }
    long timeout_ms = argc > 3 ? atol(argv[3]) : 10000;
    float vec[SPLINTER_EMBED_DIM];
    FILE *fp = fopen(argv[2], "rb");
    if (!fp || fread(vec, sizeof(float), SPLINTER_EMBED_DIM, fp) != SPLINTER_EMBED_DIM) return 3;
    fclose(fp);
    if (splinter_open(argv[1]) != 0) return 4;
    splinter_header_snapshot_t hs;
    splinter_get_header_snapshot(&hs);
    char **keys = calloc(hs.slots, sizeof(char *));
    struct timespec nap = { 0, 2 * 1000000 };
    for (long waited = 0; waited < timeout_ms; waited += 2) {
        size_t n = 0;
        splinter_list(keys, hs.slots, &n);
        for (size_t i = 0; i < n; i++) {
            if (strncmp(keys[i], "__sqtmp_", 8) != 0) continue;
            char key[SPLINTER_KEY_MAX];
            strncpy(key, keys[i], SPLINTER_KEY_MAX - 1);
            key[SPLINTER_KEY_MAX - 1] = '\0';
            splinter_slot_snapshot_t s;
            if (splinter_get_slot_snapshot(key, &s) == 0 && (s.bloom & 1ULL)) {
                struct timespec settle = { 0, 30 * 1000000 };  /* let the client enter splinter_poll */
                nanosleep(&settle, NULL);
                return splinter_set_embedding(key, vec) == 0 ? 0 : 5;
            }
        }
        nanosleep(&nap, NULL);
    }
    return 6;
}
