/*
 * splinter_gpu_search -- the reference's `search` command with the scoring loop moved to the GPU.
 *
 * Same flags, same query-vector handshake, same output as cmd_search
 * (/root/reference/splinter_cli_cmd_search.c:136-486):
 *
 *   splinter_gpu_search --use STORE <query>|- [--file PATH] [--json] [--limit N] [--distance F]
 *                       [--similarity F] [--bloom HEXMASK] [--regex PATTERN]
 *
 * What stays on the host, as in the reference: flag parsing (:186-215), reading the query text
 * (:242-275), the scratch-key handshake with the embedding daemon (:301-327), the POSIX-regex key filter
 * (3rdparty/libgrawk.c:35-37,62), printing (:419-473).  What moved: candidate enumeration, per-key
 * snapshot + embedding fetch, both scorers, thresholds, sort, limit (:336-416) -> one spl_gpu_search call
 * (SPL_MODE_EXACT, so the printed scores are the reference's bit for bit).
 *
 * The reference's store library is loaded at run time (dlopen "libsplinter.so", or $SPLINTER_LIB): it
 * stays the owner of the store format and of every mutation.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <errno.h>
#include <float.h>
#include <getopt.h>
#include <regex.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "splinter_gpu.h"

#define KEY_MAX 64
#define TYPE_VARTEXT (1u << 7) /* SPL_SLOT_TYPE_VARTEXT, splinter.h */

/* header snapshot as laid out by splinter.h:261-280 */
typedef struct {
    uint32_t magic, version, slots, max_val_sz;
    uint64_t epoch;
    uint8_t core_flags, user_flags;
    uint64_t parse_failures, last_failure_epoch;
} header_snapshot_t;

static struct {
    int (*open)(const char *);
    void (*close)(void);
    int (*set)(const char *, const void *, size_t);
    int (*unset)(const char *);
    int (*set_named_type)(const char *, uint16_t);
    int (*set_label)(const char *, uint64_t);
    int (*bump_slot)(const char *);
    int (*poll)(const char *, uint64_t);
    int (*get_embedding)(const char *, float *);
    int (*get_header_snapshot)(header_snapshot_t *);
} R;

static int load_reference(void) {
    const char *path = getenv("SPLINTER_LIB");
    void *h = dlopen(path && *path ? path : "libsplinter.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { fprintf(stderr, "search: cannot load the splinter store library: %s\n", dlerror()); return -1; }
#define SYM(field, name) do { *(void **)(&R.field) = dlsym(h, name); if (!R.field) { fprintf(stderr, "search: %s missing\n", name); return -1; } } while (0)
    SYM(open, "splinter_open"); SYM(close, "splinter_close"); SYM(set, "splinter_set"); SYM(unset, "splinter_unset");
    SYM(set_named_type, "splinter_set_named_type"); SYM(set_label, "splinter_set_label"); SYM(bump_slot, "splinter_bump_slot");
    SYM(poll, "splinter_poll"); SYM(get_embedding, "splinter_get_embedding"); SYM(get_header_snapshot, "splinter_get_header_snapshot");
#undef SYM
    return 0;
}

static const char *show_key_type(unsigned flags) { /* labels of splinter_cli_util.c:128-139 */
    static const struct { unsigned bit; const char *name; } t[] = {
        { 1u << 1, "SPL_SLOT_TYPE_BIGINT" }, { 1u << 2, "SPL_SLOT_TYPE_BIGUINT" }, { 1u << 4, "SPL_SLOT_TYPE_BINARY" },
        { 1u << 5, "SPL_SLOT_TYPE_IMGDATA" }, { 1u << 7, "SPL_SLOT_TYPE_VARTEXT" }, { 1u << 6, "SPL_SLOT_TYPE_AUDIO" },
        { 1u << 3, "SPL_SLOT_TYPE_JSON" }, { 1u << 0, "SPL_SLOT_TYPE_VOID" } };
    for (size_t i = 0; i < sizeof t / sizeof t[0]; i++) if (flags & t[i].bit) return t[i].name;
    return "UNNAMED";
}

static void usage(void) {
    printf("search searches embedded keys by semantic similarity.\n");
    printf("Usage: search <query>|- [--file PATH] [--json] [--limit N] [--distance F] [--similarity F] [--bloom MASK] [--regex PATTERN]\n");
}

static char *slurp(FILE *fp, size_t cap, size_t *out_len, int *truncated) {
    if (cap == 0) return NULL;
    char *buf = malloc(cap + 1);
    if (!buf) return NULL;
    size_t total = 0, got;
    while (total < cap && (got = fread(buf + total, 1, cap - total, fp)) > 0) total += got;
    if (total == cap) { char probe; if (fread(&probe, 1, 1, fp) == 1) { *truncated = 1; while (fread(&probe, 1, 1, fp) == 1) {} } }
    buf[total] = '\0';
    *out_len = total;
    return buf;
}

typedef struct { char key[KEY_MAX]; uint32_t val_len; uint8_t type_flag; } hit_meta_t;

int main(int argc, char **argv) {
    static const struct option opts[] = {
        { "use", required_argument, NULL, 'u' }, { "json", no_argument, NULL, 'j' }, { "limit", required_argument, NULL, 'L' },
        { "distance", required_argument, NULL, 'd' }, { "similarity", required_argument, NULL, 's' },
        { "bloom", required_argument, NULL, 'B' }, { "regex", required_argument, NULL, 'r' },
        { "file", required_argument, NULL, 'f' }, { NULL, 0, NULL, 0 } };
    const char *store = getenv("SPLINTER_DEFAULT_STORE"), *file_path = NULL, *regex_pattern = NULL;
    char *query_text = NULL, *query_buffer = NULL;
    size_t query_len = 0;
    int use_json = 0, limit = 0, read_stdin = 0, embedding_available = 1, rc = -1, scratch_written = 0;
    float max_distance = 0.0f, min_similarity = 0.0f;
    uint64_t bloom_mask = 0;
    int opt;
    while ((opt = getopt_long(argc, argv, "-u:jL:d:s:B:r:f:", opts, NULL)) != -1) {
        switch (opt) {
            case 'u': store = optarg; break;
            case 'j': use_json = 1; break;
            case 'L': limit = atoi(optarg); break;
            case 'd': max_distance = (float)atof(optarg); break;
            case 's': min_similarity = (float)atof(optarg); break;
            case 'B': bloom_mask = strtoull(optarg, NULL, 16); break;   /* hex, like _search.c:203 */
            case 'r': regex_pattern = optarg; break;
            case 'f': file_path = optarg; break;
            case 1: if (strcmp(optarg, "-") == 0) read_stdin = 1; else if (!query_text) query_text = optarg; break;
            default: break;
        }
    }
    if (file_path && (query_text || read_stdin)) { fprintf(stderr, "search: --file is mutually exclusive with a positional query.\n"); return 1; }
    if (!file_path && !query_text && !read_stdin) { usage(); return 1; }
    if (!store) { fprintf(stderr, "search: not connected to a store.\n"); return 1; }
    if (load_reference() != 0) return 1;
    if (R.open(store) != 0) { fprintf(stderr, "search: not connected to a store.\n"); return 1; }
    header_snapshot_t snap;
    memset(&snap, 0, sizeof snap);
    R.get_header_snapshot(&snap);
    if (snap.slots == 0) { fprintf(stderr, "search: no slots available in current store.\n"); return 1; }

    if (!query_text) {
        FILE *fp = (read_stdin || strcmp(file_path, "-") == 0) ? stdin : fopen(file_path, "rb");
        const char *label = fp == stdin ? "stdin" : file_path;
        if (!fp) { fprintf(stderr, "search: cannot open '%s': %s\n", file_path, strerror(errno)); return 1; }
        int truncated = 0;
        query_buffer = slurp(fp, snap.max_val_sz > 0 ? (size_t)snap.max_val_sz - 1 : 0, &query_len, &truncated);
        if (fp != stdin) fclose(fp);
        if (!query_buffer || query_len == 0) { fprintf(stderr, "search: no query text read from %s.\n", label); free(query_buffer); return 1; }
        if (truncated) fprintf(stderr, "search: warning: %s exceeded store max_val_sz (%u); truncated to %zu bytes.\n", label, snap.max_val_sz, query_len);
        query_text = query_buffer;
    }

    /* ---- query-vector handshake, exactly as _search.c:301-327 ---- */
    char scratch_key[KEY_MAX];
    snprintf(scratch_key, sizeof scratch_key, "__sqtmp_%d", (int)getpid());
    R.set(scratch_key, query_text, query_buffer ? query_len : strlen(query_text));
    R.set_named_type(scratch_key, TYPE_VARTEXT);
    R.set_label(scratch_key, 1ULL);
    R.bump_slot(scratch_key);
    scratch_written = 1;
    int poll_rc = R.poll(scratch_key, 2000);
    float *query_vec = NULL;
    spl_gpu_ctx *gpu = NULL;
    spl_gpu_result *hits = NULL;
    hit_meta_t *meta = NULL;
    regex_t rx;
    int have_rx = 0;

    gpu = spl_gpu_attach(store, NULL);
    if (!gpu) { fprintf(stderr, "search: %s\n", spl_gpu_last_error()); goto cleanup; }
    spl_gpu_info info;
    spl_gpu_get_info(gpu, &info);
    query_vec = calloc(info.dim, sizeof(float));
    if (poll_rc == 0) {
        if (R.get_embedding(scratch_key, query_vec) != 0) {
            fprintf(stderr, "search: warning: could not retrieve query embedding; scoring unavailable.\n");
            embedding_available = 0;
        }
    } else if (poll_rc == -1 && errno == ETIMEDOUT) {
        fprintf(stderr, "search: warning: splinference timed out; embedding scoring unavailable.\n");
        embedding_available = 0;
    } else if (poll_rc == -2) {
        fprintf(stderr, "search: store error during poll.\n");
        goto cleanup;
    }
    if (!embedding_available && !regex_pattern && bloom_mask == 0) {
        fprintf(stderr, "search: splinference unavailable and no regex/bloom filter; no results.\n");
        rc = 0;
        goto cleanup;
    }
    if (regex_pattern) {
        if (regcomp(&rx, regex_pattern, REG_NOSUB | REG_NEWLINE | REG_EXTENDED) != 0) { fprintf(stderr, "search: bad regex.\n"); goto cleanup; }
        have_rx = 1;
    }

    /* the scratch key's slot: val_off = slot * max_val_sz (splinter.c:228), found through the GPU lib's key lookup */
    uint32_t scratch_slot = SPL_NO_SLOT;
    for (uint32_t i = 0; i < info.total_slots; i++) {
        char k[KEY_MAX]; uint32_t vl; uint8_t tf;
        if (spl_gpu_slot_meta(gpu, i, k, &vl, &tf) == 0 && strncmp(k, scratch_key, KEY_MAX) == 0) { scratch_slot = i; break; }
    }
    if (!embedding_available) memset(query_vec, 0, info.dim * sizeof(float));  /* nothing to score against */

    /* ---- the hot path: one call ---- */
    spl_gpu_search_params p;
    memset(&p, 0, sizeof p);
    p.struct_size = sizeof p; p.nq = 1; p.dim = info.dim;
    /* the regex is applied before the limit in the reference (:365-368 then :415-416): ask for everything then */
    /* ... and "active_keys" counts every survivor; the fused top-k path reports that count exactly unless a
     * threshold flag is active (then the sorted full pass is used). */
    const int thresholds = min_similarity > 0.0f || max_distance > 0.0f;
    p.k = (limit > 0 && !have_rx && embedding_available && !thresholds) ? (uint32_t)limit : 0;
    p.order = SPL_ORDER_COSINE; p.mode = SPL_MODE_EXACT; p.bloom_mask = bloom_mask;
    p.min_similarity = embedding_available ? min_similarity : 0.f;
    p.max_distance = embedding_available ? max_distance : 0.f;
    p.exclude_slot = scratch_slot;
    uint32_t cap = p.k ? p.k : (uint32_t)info.total_slots, n_hits = 0;
    hits = calloc(cap, sizeof *hits);
    if (spl_gpu_search(gpu, &p, query_vec, hits, cap, &n_hits) != 0) { fprintf(stderr, "search: %s\n", spl_gpu_last_error()); goto cleanup; }
    if (n_hits > cap) n_hits = cap;
    uint32_t total_survivors = n_hits;
    if (p.k) spl_gpu_last_totals(gpu, &total_survivors, 1);

    /* resolve keys, apply the regex, then the limit */
    meta = calloc(n_hits ? n_hits : 1, sizeof *meta);
    uint32_t result_count = 0;
    for (uint32_t i = 0; i < n_hits; i++) {
        hit_meta_t m;
        if (spl_gpu_slot_meta(gpu, hits[i].slot_idx, m.key, &m.val_len, &m.type_flag) != 0) continue;
        if (have_rx && regexec(&rx, m.key, 0, NULL, 0) != 0) continue;
        if (!embedding_available) {
            if (min_similarity > 0.0f || max_distance > 0.0f) continue;     /* _search.c:405 */
            hits[i].has_embedding = 0;
        }
        hits[result_count] = hits[i];
        meta[result_count++] = m;
    }
    if (p.k) result_count = total_survivors;   /* nothing was filtered on the host in this case */
    uint32_t print_count = p.k ? n_hits : result_count;
    if (limit > 0 && print_count > (uint32_t)limit) print_count = (uint32_t)limit;

    /* ---- output: format strings of _search.c:419-473 ---- */
    if (use_json) {
        printf("{\n  \"query\": \"%s\",\n  \"store\": {\n    \"total_slots\": %u,\n    \"active_keys\": %d\n  },\n  \"results\": [\n",
               query_text, snap.slots, (int)result_count);
        for (uint32_t i = 0; i < print_count; i++) {
            printf("    {\n      \"key\": \"%s\",\n", meta[i].key);
            if (hits[i].has_embedding) printf("      \"similarity\": %.4f,\n      \"distance\": %.4f,\n", hits[i].similarity, hits[i].distance);
            else printf("      \"similarity\": null,\n      \"distance\": null,\n");
            printf("      \"epoch\": %lu,\n      \"value_length\": %u,\n      \"type\": \"%s\",\n      \"bloom\": \"0x%016llx\",\n      \"has_embedding\": %s\n",
                   (unsigned long)hits[i].epoch, meta[i].val_len, show_key_type(meta[i].type_flag),
                   (unsigned long long)hits[i].bloom, hits[i].has_embedding ? "true" : "false");
            printf(i + 1 < print_count ? "    },\n" : "    }\n");
        }
        printf("  ]\n}\n");
    } else {
        printf("%-44s %-10s %-10s %-6s %-6s %s\n", "Name", "Similarity", "Distance", "Epoch", "Len", "Type");
        for (uint32_t i = 0; i < print_count; i++) {
            if (hits[i].has_embedding)
                printf("%-44s %-10.4f %-10.4f %-6lu %-6u %s\n", meta[i].key, hits[i].similarity, hits[i].distance,
                       (unsigned long)hits[i].epoch, meta[i].val_len, show_key_type(meta[i].type_flag));
            else
                printf("%-44s %-10s %-10s %-6lu %-6u %s\n", meta[i].key, "-", "-", (unsigned long)hits[i].epoch, meta[i].val_len,
                       show_key_type(meta[i].type_flag));
        }
        puts("");
    }
    rc = 0;

cleanup:
    if (scratch_written) R.unset(scratch_key);
    if (have_rx) regfree(&rx);
    if (gpu) spl_gpu_detach(gpu);
    free(hits); free(meta); free(query_vec); free(query_buffer);
    return rc == 0 ? 0 : 1;
}
