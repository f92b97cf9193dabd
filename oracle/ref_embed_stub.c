/*
 * oracle/ref_embed_stub.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Stands in for the `splinference` daemon (llama.cpp; not buildable here) for ANY search client: it
 * watches the store for a scratch key `__sqtmp_<pid>` carrying label 0x1 (the handshake of
 * /root/reference/splinter_cli_cmd_search.c:301-311) and answers it with splinter_set_embedding(vector
 * read from a file), through the reference's own API.
 *
 *   ref_embed_stub <store> <query.f32> [timeout_ms] [count]
 *
 * count (default 1) = how many handshakes to answer before exiting (a REPL session asks several times under
 * the same scratch key: a handshake is "open" while the key's embedding is still all zero).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "splinter.h"

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s <store> <query.f32> [timeout_ms]\n", argv[0]); return 2; }
    long timeout_ms = argc > 3 ? atol(argv[3]) : 10000;
    long count = argc > 4 ? atol(argv[4]) : 1, served = 0;

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
                /* a fresh handshake sits on a slot whose embedding splinter_set zeroed (splinter.c:406-411); one that
                 * still carries a vector was answered already and has not been unset by its client yet */
                int answered = 0;
                for (int d = 0; d < SPLINTER_EMBED_DIM && !answered; d++) answered = s.embedding[d] != 0.0f;
                if (answered) continue;
                struct timespec settle = { 0, 30 * 1000000 };  /* let the client enter splinter_poll */
                nanosleep(&settle, NULL);
                if (splinter_set_embedding(key, vec) != 0) return 5;
                if (++served >= count) return 0;
            }
        }
        nanosleep(&nap, NULL);
    }
    return 6;
}
