/*
 * splinter_cli_cmd_search_gpu.c -- the reference CLI's `search` module with its scoring loop on the GPU.
 *
 * Drop-in for /root/reference/splinter_cli_cmd_search.c: it defines the same two module entry points the
 * reference registers in command_modules[] (splinter_cli_main.c:268-277; types splinter_cli.h:19-21,
 * prototypes :165-166)
 *
 *     int  cmd_search(int argc, char *argv[]);
 *     void help_cmd_search(unsigned int level);
 *
 * and is compiled against the reference's own splinter_cli.h / splinter.h / 3rdparty/grawk.h.  Build the
 * reference CLI with this file in place of splinter_cli_cmd_search.c and link -lsplinter_gpu: nothing else
 * in the reference changes (oracle/Makefile builds exactly that as _ref/gpu/splinterctl for the tests).
 *
 * What stays on the host, with the reference's behaviour: flag parsing (_search.c:170-215), reading the
 * query text (:242-275), the scratch-key handshake with the embedding daemon (:301-327), the key regex
 * (grawk, :349-368), printing (:419-473), return codes.  What moved to libsplinter_gpu.so (:329-416):
 * candidate enumeration, per-key snapshot, embedding fetch, both scorers, thresholds, sort, limit -- one
 * spl_gpu_search call in SPL_MODE_EXACT, so the printed scores are the reference's bit for bit.
 *
 * The GPU context is attached once per process and store and kept: in the REPL the first `search` pays
 * for the CUDA context and the full upload of the store, every later one only for an incremental
 * spl_gpu_sync of the slots that changed.
 */
#ifdef HAVE_EMBEDDINGS
#include <errno.h>
#include <float.h>
#include <getopt.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "splinter_cli.h"
#include "3rdparty/grawk.h"
#include "splinter_gpu.h"

static const char *modname = "search";

/* ---- one mirror per process, re-attached when the user switches stores (`use`) ---------------------- */
static spl_gpu_ctx *g_gpu = NULL;
static char g_gpu_store[sizeof(((cli_user_t *)0)->store)];
static int g_gpu_atexit = 0;

static void gpu_release(void) {
    if (g_gpu) spl_gpu_detach(g_gpu);
    g_gpu = NULL;
    g_gpu_store[0] = '\0';
}

static spl_gpu_ctx *gpu_for_current_store(void) {
    if (g_gpu && strncmp(g_gpu_store, thisuser.store, sizeof g_gpu_store) == 0) {
        if (spl_gpu_sync(g_gpu, NULL) == 0) return g_gpu;
        gpu_release();   /* the store went away or was re-created under us: attach afresh */
    } else if (g_gpu) {
        gpu_release();
    }
    spl_gpu_opts o;
    memset(&o, 0, sizeof o);
    o.struct_size = sizeof o;
    o.device = -1;
#ifdef SPLINTER_PERSISTENT
    o.flags |= SPL_OPT_PERSISTENT;   /* the reference's file-backed build opens a path (splinter.c:237-241) */
#endif
    g_gpu = spl_gpu_attach(thisuser.store, &o);   /* includes the first full spl_gpu_sync */
    if (!g_gpu) return NULL;
    snprintf(g_gpu_store, sizeof g_gpu_store, "%s", thisuser.store);
    if (!g_gpu_atexit) { atexit(gpu_release); g_gpu_atexit = 1; }
    return g_gpu;
}

void help_cmd_search(unsigned int level) {
    (void)level;
    printf("%s searches embedded keys by semantic similarity.\n", modname);
    printf("Usage: %s <query>|- [--file PATH] [--json] [--limit N] [--distance F] [--similarity F] [--bloom MASK] [--regex PATTERN]\n", modname);
    printf("  <query>         Quoted query string. Use '-' to read the query from stdin.\n");
    printf("  --file PATH     Read query text from PATH (use '-' for stdin). Mutually exclusive with positional query.\n");
    printf("  --json          Output results as JSON\n");
    printf("  --limit N       Maximum number of results (default: unlimited)\n");
    printf("  --distance F    Maximum euclidean distance to include (default: no filter)\n");
    printf("  --similarity F  Minimum cosine similarity to include (default: no filter)\n");
    printf("  --bloom MASK    Hex bloom mask; uses enumerate_matches() for pre-filtering\n");
    printf("  --regex PATTERN Regex filter applied to key names before scoring\n");
}

static const struct option gpu_search_options[] = {
    { "json", no_argument, NULL, 'j' },           { "limit", required_argument, NULL, 'L' },
    { "distance", required_argument, NULL, 'd' }, { "similarity", required_argument, NULL, 's' },
    { "bloom", required_argument, NULL, 'B' },    { "regex", required_argument, NULL, 'r' },
    { "file", required_argument, NULL, 'f' },     { NULL, 0, NULL, 0 }
};

/* read at most cap bytes; tell the caller whether the source held more (and drain it) */
static char *read_query_stream(FILE *fp, size_t cap, size_t *len, int *truncated) {
    if (cap == 0) return NULL;
    char *buf = (char *)malloc(cap + 1);
    if (!buf) return NULL;
    size_t n = 0, got;
    while (n < cap && (got = fread(buf + n, 1, cap - n, fp)) > 0) n += got;
    if (n == cap) {
        char extra;
        if (fread(&extra, 1, 1, fp) == 1) {
            *truncated = 1;
            while (fread(&extra, 1, 1, fp) == 1) {}
        }
    }
    buf[n] = '\0';
    *len = n;
    return buf;
}

typedef struct { char key[SPLINTER_KEY_MAX]; uint32_t val_len; uint8_t type_flag; } hit_meta_t;

int cmd_search(int argc, char *argv[]) {
    const char *file_path = NULL;
    char *query_text = NULL, *query_buffer = NULL, *regex_pattern = NULL;
    size_t query_len = 0;
    int use_json = 0, limit = 0, stdin_positional = 0, embedding_available = 1, scratch_written = 0, rc = -1;
    float max_distance = 0.0f, min_similarity = 0.0f;
    uint64_t bloom_mask = 0;
    grawk_t *g = NULL;
    awk_pat_t *filter = NULL;
    spl_gpu_result *hits = NULL;
    hit_meta_t *meta = NULL;
    splinter_header_snapshot_t snap;
    char scratch_key[SPLINTER_KEY_MAX];
    snprintf(scratch_key, sizeof scratch_key, "__sqtmp_%d", (int)getpid());

    if (argc < 2) { help_cmd_search(1); return -1; }
    /* argv[1] is the query, "-" (stdin) or already a flag (_search.c:170-181) */
    if (strcmp(argv[1], "-") == 0) { stdin_positional = 1; optind = 2; }
    else if (argv[1][0] != '-') { query_text = argv[1]; optind = 2; }
    else optind = 1;
    int opt;
    while ((opt = getopt_long(argc, argv, "jL:d:s:B:r:f:", gpu_search_options, NULL)) != -1) {
        switch (opt) {
            case 'j': use_json = 1; break;
            case 'L': limit = atoi(optarg); break;
            case 'd': max_distance = (float)atof(optarg); break;
            case 's': min_similarity = (float)atof(optarg); break;
            case 'B': bloom_mask = strtoull(optarg, NULL, 16); break;   /* hex, _search.c:203 */
            case 'r': regex_pattern = optarg; break;
            case 'f': file_path = optarg; break;
            default: break;
        }
    }
    if (file_path && (query_text || stdin_positional)) {
        fprintf(stderr, "%s: --file is mutually exclusive with a positional query.\n", modname);
        return -1;
    }
    if (!file_path && !query_text && !stdin_positional) { help_cmd_search(1); return -1; }
    if (!thisuser.store_conn) { fprintf(stderr, "%s: not connected to a store.\n", modname); return -1; }
    memset(&snap, 0, sizeof snap);
    splinter_get_header_snapshot(&snap);
    if (snap.slots == 0) { fprintf(stderr, "%s: no slots available in current store.\n", modname); return -1; }

    if (!query_text) {
        const int from_stdin = stdin_positional || strcmp(file_path, "-") == 0;
        FILE *fp = from_stdin ? stdin : fopen(file_path, "rb");
        const char *label = from_stdin ? "stdin" : file_path;
        if (!fp) { fprintf(stderr, "%s: cannot open '%s': %s\n", modname, file_path, strerror(errno)); return -1; }
        int truncated = 0;
        query_buffer = read_query_stream(fp, snap.max_val_sz > 0 ? (size_t)snap.max_val_sz - 1 : 0, &query_len, &truncated);
        if (!from_stdin) fclose(fp);
        if (!query_buffer || query_len == 0) {
            fprintf(stderr, "%s: no query text read from %s.\n", modname, label);
            free(query_buffer);
            return -1;
        }
        if (truncated)
            fprintf(stderr, "%s: warning: %s exceeded store max_val_sz (%u); truncated to %zu bytes.\n", modname, label,
                    snap.max_val_sz, query_len);
        query_text = query_buffer;
    }

    /* Bring the mirror up to date BEFORE the handshake: splinter_poll only sees an answer that arrives after it
     * was called (splinter.c:479-516), so nothing slow -- a first attach creates the CUDA context and uploads the
     * whole store -- may sit between the bump below and the poll. */
    spl_gpu_ctx *gpu = gpu_for_current_store();

    /* ---- query-vector handshake with the embedding daemon (_search.c:301-327), unchanged ---- */
    splinter_set(scratch_key, query_text, query_buffer ? query_len : strlen(query_text));
    splinter_set_named_type(scratch_key, SPL_SLOT_TYPE_VARTEXT);
    splinter_set_label(scratch_key, 1ULL);
    splinter_bump_slot(scratch_key);
    scratch_written = 1;
    const int poll_rc = splinter_poll(scratch_key, 2000);
    float query_vec[SPLINTER_EMBED_DIM];
    memset(query_vec, 0, sizeof query_vec);
    if (poll_rc == 0) {
        if (splinter_get_embedding(scratch_key, query_vec) != 0) {
            fprintf(stderr, "%s: warning: could not retrieve query embedding; scoring unavailable.\n", modname);
            embedding_available = 0;
        }
    } else if (poll_rc == -1 && errno == ETIMEDOUT) {
        fprintf(stderr, "%s: warning: splinference timed out; embedding scoring unavailable.\n", modname);
        embedding_available = 0;
    } else if (poll_rc == -2) {
        fprintf(stderr, "%s: store error during poll.\n", modname);
        goto cleanup;
    }
    if (!embedding_available && !regex_pattern && bloom_mask == 0) {
        fprintf(stderr, "%s: splinference unavailable and no regex/bloom filter; no results.\n", modname);
        rc = 0;
        goto cleanup;
    }
    if (!gpu) { fprintf(stderr, "%s: %s\n", modname, spl_gpu_last_error()); goto cleanup; }
    spl_gpu_info info;
    spl_gpu_get_info(gpu, &info);
    if (info.dim != SPLINTER_EMBED_DIM) {
        fprintf(stderr, "%s: store holds %u-d embeddings, this build %d-d.\n", modname, info.dim, SPLINTER_EMBED_DIM);
        goto cleanup;
    }
    if (regex_pattern) {
        static const grawk_opts_t gopts = { 0, 0, 1 };   /* invert_match, ignore_case, quiet (_search.c:139-143) */
        g = grawk_init();
        if (!g) { fprintf(stderr, "%s: unable to allocate memory to filter keys.\n", modname); errno = ENOMEM; goto cleanup; }
        grawk_set_options(g, &gopts);
        filter = grawk_build_pattern(regex_pattern);
        grawk_set_pattern(g, filter);
    }

    /* ---- the hot path (_search.c:329-416): one call ---- */
    spl_gpu_search_params p;
    memset(&p, 0, sizeof p);
    p.struct_size = sizeof p;
    p.nq = 1;
    p.dim = info.dim;
    /* The regex runs before the limit in the reference (:365-368, then :415-416), and "active_keys" counts every
     * survivor: ask for everything when a regex or a threshold is active; otherwise the fused top-k reports the
     * survivor count on the side. */
    const int thresholds = min_similarity > 0.0f || max_distance > 0.0f;
    p.k = (limit > 0 && !filter && embedding_available && !thresholds) ? (uint32_t)limit : 0;
    p.order = SPL_ORDER_COSINE;
    p.mode = SPL_MODE_EXACT;
    p.bloom_mask = bloom_mask;
    p.min_similarity = embedding_available ? min_similarity : 0.f;
    p.max_distance = embedding_available ? max_distance : 0.f;
    p.exclude_slot = SPL_NO_SLOT;
    spl_gpu_find_key(gpu, scratch_key, &p.exclude_slot);   /* the scratch key never appears in the results (:367) */
    if (spl_gpu_sync(gpu, NULL) != 0) { fprintf(stderr, "%s: %s\n", modname, spl_gpu_last_error()); goto cleanup; }
    uint32_t cap = p.k ? p.k : (uint32_t)info.total_slots, n_hits = 0;
    hits = (spl_gpu_result *)calloc(cap, sizeof *hits);
    if (!hits) { fprintf(stderr, "%s: unable to allocate memory for results.\n", modname); errno = ENOMEM; goto cleanup; }
    if (spl_gpu_search(gpu, &p, query_vec, hits, cap, &n_hits) != 0) {
        fprintf(stderr, "%s: %s\n", modname, spl_gpu_last_error());
        goto cleanup;
    }
    if (n_hits > cap) n_hits = cap;
    uint32_t survivors = n_hits;
    if (p.k) spl_gpu_last_totals(gpu, &survivors, 1);

    /* keys / lengths / types of the hits, then the regex, then the limit */
    meta = (hit_meta_t *)calloc(n_hits ? n_hits : 1, sizeof *meta);
    if (!meta) { fprintf(stderr, "%s: unable to allocate memory for results.\n", modname); errno = ENOMEM; goto cleanup; }
    uint32_t result_count = 0;
    for (uint32_t i = 0; i < n_hits; i++) {
        hit_meta_t m;
        if (spl_gpu_slot_meta(gpu, hits[i].slot_idx, m.key, &m.val_len, &m.type_flag) != 0) continue;
        if (filter && !grawk_match(g, m.key)) continue;
        if (!embedding_available) {
            if (thresholds) continue;   /* _search.c:405 */
            hits[i].has_embedding = 0;
        }
        hits[result_count] = hits[i];
        meta[result_count++] = m;
    }
    if (p.k) result_count = survivors;   /* nothing was filtered on the host in that case */
    int print_count = (int)(p.k ? n_hits : result_count);
    if (limit > 0 && print_count > limit) print_count = limit;

    /* ---- output, in the reference's format (_search.c:419-473) ---- */
    if (use_json) {
        printf("{\n  \"query\": \"%s\",\n  \"store\": {\n    \"total_slots\": %u,\n    \"active_keys\": %d\n  },\n  \"results\": [\n",
               query_text, snap.slots, (int)result_count);
        for (int i = 0; i < print_count; i++) {
            printf("    {\n      \"key\": \"%s\",\n", meta[i].key);
            if (hits[i].has_embedding)
                printf("      \"similarity\": %.4f,\n      \"distance\": %.4f,\n", hits[i].similarity, hits[i].distance);
            else
                printf("      \"similarity\": null,\n      \"distance\": null,\n");
            printf("      \"epoch\": %lu,\n      \"value_length\": %u,\n      \"type\": \"%s\",\n", (unsigned long)hits[i].epoch,
                   meta[i].val_len, cli_show_key_type((unsigned short)meta[i].type_flag));
            printf("      \"bloom\": \"0x%016llx\",\n      \"has_embedding\": %s\n", (unsigned long long)hits[i].bloom,
                   hits[i].has_embedding ? "true" : "false");
            printf(i < print_count - 1 ? "    },\n" : "    }\n");
        }
        printf("  ]\n}\n");
    } else {
        printf("%-44s %-10s %-10s %-6s %-6s %s\n", "Name", "Similarity", "Distance", "Epoch", "Len", "Type");
        for (int i = 0; i < print_count; i++) {
            if (hits[i].has_embedding)
                printf("%-44s %-10.4f %-10.4f %-6lu %-6u %s\n", meta[i].key, hits[i].similarity, hits[i].distance,
                       (unsigned long)hits[i].epoch, meta[i].val_len, cli_show_key_type((unsigned short)meta[i].type_flag));
            else
                printf("%-44s %-10s %-10s %-6lu %-6u %s\n", meta[i].key, "-", "-", (unsigned long)hits[i].epoch,
                       meta[i].val_len, cli_show_key_type((unsigned short)meta[i].type_flag));
        }
        puts("");
    }
    rc = 0;

cleanup:
    if (scratch_written) splinter_unset(scratch_key);
    if (g) grawk_free(g);
    free(hits);
    free(meta);
    free(query_buffer);
    return rc;
}
#endif /* HAVE_EMBEDDINGS */
