/*
 * oracle/ref_search_cli.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Runs the reference's real `search` command (cmd_search,
 * /root/reference/splinter_cli_cmd_search.c:136-486) end to end against a live
 * store.  The query embedding normally comes from the `splinference` daemon
 * (llama.cpp; not buildable here) through the scratch-key handshake at
 * _search.c:301-327.  A forked child stands in for that daemon: it waits for
 * the parent's scratch key `__sqtmp_<pid>` to carry label 0x1 and then answers
 * with splinter_set_embedding(<vector read from a file>).
 *
 *   ref_search_cli <store> <query.f32> <search args...>
 *     e.g. ref_search_cli /spl_t1 q.bin "some text" --limit 10 --bloom 3 --json
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
#include <sys/wait.h>

#include "splinter.h"

extern int ref_cmd_search(int argc, char **argv);

static void stub_embedder(const char *store, const char *vec_path, pid_t parent) {
    float vec[SPLINTER_EMBED_DIM];
    FILE *fp = fopen(vec_path, "rb");
    if (!fp || fread(vec, sizeof(float), SPLINTER_EMBED_DIM, fp) != SPLINTER_EMBED_DIM) _exit(3);
    fclose(fp);
    if (splinter_open(store) != 0) _exit(4);
    char key[SPLINTER_KEY_MAX];
    snprintf(key, sizeof key, "__sqtmp_%d", (int)parent);
    struct timespec nap = { 0, 1000000 };
    for (int spins = 0; spins < 4000; spins++) {
        splinter_slot_snapshot_t s;
        /* label 0x1 is set just before the bump that precedes splinter_poll (:307-311) */
        if (splinter_get_slot_snapshot(key, &s) == 0 && (s.bloom & 1ULL)) {
            struct timespec settle = { 0, 30 * 1000000 }; /* let the parent enter splinter_poll */
            nanosleep(&settle, NULL);
            _exit(splinter_set_embedding(key, vec) == 0 ? 0 : 5);
        }
        nanosleep(&nap, NULL);
    }
    _exit(6);
}

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s <store> <query.f32> <search args...>\n", argv[0]);
        return 2;
    }
    pid_t parent = getpid();
    pid_t child = fork();
    if (child == 0) stub_embedder(argv[1], argv[2], parent);
    if (splinter_open(argv[1]) != 0) { perror("splinter_open"); return 1; }
    /* argv for cmd_search: argv[0]="search", then the user's args */
    int sargc = argc - 2;
    char **sargv = (char **)calloc((size_t)sargc + 1, sizeof(char *));
    sargv[0] = "search";
    for (int i = 1; i < sargc; i++) sargv[i] = argv[i + 2];
    int rc = ref_cmd_search(sargc, sargv);
    fflush(stdout);
    int st = 0;
    waitpid(child, &st, 0);
    free(sargv);
    return rc == 0 ? 0 : 1;
}
