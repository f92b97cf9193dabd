/*
 * splinter_gpu.h -- C ABI of libsplinter_gpu.so: the B200 (sm_100a) drop-in for
 * libsplinter's one data-parallel hot path, the brute-force similarity / distance
 * scan over per-slot embeddings with the 64-bit bloom-label pre-filter.
 *
 * The reference has no library-level search call: the whole path lives in its CLI
 * command (splinter_cli_cmd_search.c:136-486) on top of five store calls.  This
 * header therefore DEFINES the search entry points, with the reference's semantics:
 *
 *   reference (file:line)                                  replaced by
 *   ------------------------------------------------------ -------------------------
 *   splinter_open + static H/S/VALUES (splinter.c:46-57,   spl_gpu_attach / _attach_image
 *     235-248), magic/version gate (:246)
 *   splinter_get_slot_snapshot seqlock copy (:532-564),     spl_gpu_sync
 *     splinter_get_embedding (:590-609), dirty tracking of
 *     the writers in splinter.c:326-429,567-588,799-867
 *   splinter_enumerate_matches (:977-989), splinter_list    spl_gpu_search (bloom_mask)
 *     (:466-477)
 *   cosine_similarity / euclidean_distance /                spl_gpu_search
 *     compare_results + qsort + limit + thresholds
 *     (splinter_cli_cmd_search.c:43-72,373-416)
 *   search_result_t (_search.c:26-35)                       spl_gpu_result
 *
 * Conventions follow splinter.h:398-412: 0 = ok, -1 = recoverable (errno set:
 * EAGAIN, ENOMEM, EIO for CUDA failures, ...), -2 = caller error (errno = EINVAL).
 * spl_gpu_last_error() returns a thread-local message for the last failure.
 *
 * There is NO CPU fallback: every entry point fails loudly when no sm_100 device
 * (or no CUDA driver) is present.
 */
#ifndef SPLINTER_GPU_H
#define SPLINTER_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPL_GPU_ABI_VERSION 2

typedef struct spl_gpu_ctx spl_gpu_ctx;

/* ---- result ordering -------------------------------------------------------------- */
#define SPL_ORDER_COSINE 0u /* reference: similarity desc, then distance asc (_search.c:64-72);
                               remaining ties (unspecified by the reference's qsort) -> slot idx asc */
#define SPL_ORDER_L2     1u /* distance asc, similarity desc, idx asc (batched L2 config) */

/* ---- compute mode ------------------------------------------------------------------ */
#define SPL_MODE_AUTO    0u /* FAST (small batches share passes 8, 4 or 2 queries at a time); BATCH_BF16 from 16 queries up
                               (from 4 when no label mask is given) if a bf16 mirror exists and no threshold is set on
                               the score that does not order the result */
#define SPL_MODE_FAST    1u /* fp32, per-lane partial sums + warp-shuffle tree (|d sim| ~1e-6)           */
#define SPL_MODE_EXACT   2u /* fp32, the reference's sequential summation order: scores and order bit-identical
                               to the reference (fast candidate scan + sequential re-score with a selection proof) */
#define SPL_MODE_BATCH_BF16 3u /* tcgen05 Q x E^T on the bf16 mirror, fp32 accumulate, fp32 re-rank; thresholds only on
                                  the ordering score (cosine: min_similarity, L2: max_distance), else ENOSYS */

/* ---- attach options ---------------------------------------------------------------- */
#define SPL_OPT_PERSISTENT   0x1u /* name is a file path (open), not a shm name (shm_open): splinter.c:237-241 */
#define SPL_OPT_BF16_MIRROR  0x2u /* also keep a bf16 copy + fp32 row norms for SPL_MODE_BATCH_BF16            */

typedef struct spl_gpu_opts {
    uint32_t struct_size;   /* = sizeof(spl_gpu_opts) */
    int32_t  device;        /* CUDA ordinal; -1 = current device */
    uint32_t flags;         /* SPL_OPT_* */
    uint32_t slot_begin;    /* shard: first slot mirrored by this context */
    uint32_t slot_count;    /* shard: number of slots (0 = to the end of the store) */
    uint32_t sync_threads;  /* host threads for the snapshot sweep (0 = auto) */
} spl_gpu_opts;

/* ---- search parameters -------------------------------------------------------------- */
#define SPL_NO_SLOT 0xFFFFFFFFu

typedef struct spl_gpu_search_params {
    uint32_t struct_size;    /* = sizeof(spl_gpu_search_params) */
    uint32_t nq;             /* number of query vectors */
    uint32_t dim;            /* must equal the mirror's dimension */
    uint32_t k;              /* --limit; 0 = unlimited like the reference's default (_search.c:415-416) */
    uint32_t order;          /* SPL_ORDER_* */
    uint32_t mode;           /* SPL_MODE_* */
    uint64_t bloom_mask;     /* --bloom; 0 -> splinter_list candidates, else enumerate_matches (_search.c:337-347) */
    float    min_similarity; /* --similarity; <= 0 disables (_search.c:398) */
    float    max_distance;   /* --distance;   <= 0 disables (_search.c:399) */
    uint32_t exclude_slot;   /* slot of the caller's scratch key (_search.c:367); SPL_NO_SLOT = none.
                                Index is GLOBAL (store slot index), like every index in this API. */
    uint32_t flags;          /* SPL_SEARCH_* */
} spl_gpu_search_params;

/* spl_gpu_search_device only: this search directly follows another search of the SAME context on the SAME stream (a
 * query loop), and whatever else the caller may have enqueued on that stream in between neither writes this search's
 * query buffer nor was itself launched to overlap its predecessor.  The scan may then begin while the previous scan
 * is still merging its per-SM lists (a few microseconds during which all but one SM would idle).  Results are
 * unchanged and still land in stream order -- consecutive searches may share their output buffers.  Without the flag
 * every search waits for everything before it on the stream, as any kernel does.  (Within one call the queries of a
 * batch always overlap this way.) */
#define SPL_SEARCH_PIPELINED 0x1u

/* One hit.  Mirrors search_result_t (_search.c:26-35) with the key replaced by its slot index
 * (slot_ptr - S); key / val_len / type are resolved from the store by the host caller. */
typedef struct spl_gpu_result {
    uint32_t slot_idx;       /* global slot index */
    float    similarity;
    float    distance;       /* FLT_MAX when has_embedding == 0 (_search.c:388-391) */
    uint32_t has_embedding;
    uint64_t epoch;          /* slot epoch of the mirrored snapshot */
    uint64_t bloom;
} spl_gpu_result;

typedef struct spl_gpu_sync_stats {
    uint64_t generation;     /* snapshot id after this sync */
    uint64_t slots_swept;
    uint64_t rows_uploaded;  /* embedding rows re-sent (dirty {hash,epoch} or first sync) */
    uint64_t meta_updated;   /* bloom / validity-only updates (set_label moves no slot epoch) */
    uint64_t skipped_busy;   /* odd epoch or torn copy: previous mirror row kept, retried next sync */
    uint64_t aba_rechecked;  /* rows whose {hash, epoch} were unchanged but whose content fingerprint moved
                                (retrain's backward epoch makes the epoch alone an unreliable witness) */
    uint64_t bytes_h2d;
    double   sweep_ms, upload_ms;
    uint64_t header_epoch;   /* H->epoch observed at the start of the sweep */
    uint64_t rows_revalidated; /* clean rows whose whole content was re-hashed against the store this sweep (1/1024 of
                                  them per sweep, so every row is looked at in full once per 1024 sweeps) */
    uint64_t stale_rows_found; /* ... of which differed although {hash, epoch, first cache line} had not moved: re-sent */
} spl_gpu_sync_stats;

/* Diagnostics accumulated over the life of a context (spl_gpu_get_counters). */
typedef struct spl_gpu_counters {
    uint32_t struct_size;    /* = sizeof(spl_gpu_counters), set by the caller */
    uint32_t reserved;
    uint64_t batch_queries;  /* queries handed to the bf16 tensor-core path (host- and device-buffer calls) */
    uint64_t batch_unproven; /* ... of which the host-buffer call re-answered on the fp32 scan because their candidate
                                list overflowed or the bf16 scores could not prove the selection */
    uint64_t exact_unproven; /* SPL_MODE_EXACT queries that needed the full exact pass */
    uint64_t stale_rows_found; /* sum of spl_gpu_sync_stats.stale_rows_found */
    uint64_t batch_rescored_extra; /* rows the batched path re-scored beyond its fixed 2k per query because their bf16
                                      score was within the error bound of the k-th (host-buffer calls only) */
} spl_gpu_counters;

typedef struct spl_gpu_info {
    uint32_t abi_version, dim;
    uint64_t total_slots;    /* store geometry (H->slots) or dense rows */
    uint64_t slot_begin, slot_count;
    uint64_t generation;
    int32_t  device, sm_count;
    uint32_t has_bf16_mirror, is_dense;
    uint64_t hbm_bytes;      /* device memory held by this context */
} spl_gpu_info;

/* ---- lifecycle ---------------------------------------------------------------------- */

/* Map the store `name_or_path` read-only (same shm name / file the reference opens:
 * splinter.c:235-248), check magic/version (:246) and that the per-slot size equals
 * 128 + 4*D for some D (SURVEY 8(b) geometry gate), allocate the HBM mirror, and run a
 * first full spl_gpu_sync.  NULL on failure (errno + spl_gpu_last_error). */
spl_gpu_ctx *spl_gpu_attach(const char *name_or_path, const spl_gpu_opts *opts);

/* Same over a store image the caller has already mapped (e.g. the mapping the reference
 * library itself made).  The memory must stay mapped until spl_gpu_detach. */
spl_gpu_ctx *spl_gpu_attach_image(const void *base, size_t size, const spl_gpu_opts *opts);

/* A device-only shard with no backing store: n_rows x dim fp32 rows plus bloom / flags /
 * epoch columns, filled by spl_gpu_dense_upload or spl_gpu_dense_fill_synthetic.  Used for
 * geometries one store cannot express (D != 768, > 4 GiB arenas: splinter.h:39,194,235). */
spl_gpu_ctx *spl_gpu_open_dense(uint64_t n_rows, uint32_t dim, uint64_t global_row0, const spl_gpu_opts *opts);

void spl_gpu_detach(spl_gpu_ctx *ctx);

int  spl_gpu_get_info(const spl_gpu_ctx *ctx, spl_gpu_info *out);
int  spl_gpu_get_counters(const spl_gpu_ctx *ctx, spl_gpu_counters *out);
const char *spl_gpu_last_error(void);

/* ---- mirror maintenance ------------------------------------------------------------- */

/* Sweep the store and re-upload only dirty slots.  A slot is copied only while its seqlock
 * epoch is even and unchanged across the copy (reader rule, splinter.c:598-605); a slot that
 * is busy keeps its previous mirror row and is counted in skipped_busy.  Dirty = cached
 * {hash, epoch, bloom, val_len>0} differs from the slot's (label changes move no slot epoch:
 * splinter.c:843).
 *
 * May run concurrently with spl_gpu_search AND spl_gpu_search_device from other threads, on any stream.  What a
 * search then sees is the device-side form of the seqlock promise (splinter.h:368-396): every slot is seen either
 * as it was before or as it is after an update batch -- its embedding row, bf16 shadow, norm, flags, bloom and
 * epoch always belong to ONE consistent copy of the slot; never a half-written row.  (Different slots may be seen
 * at different points of the sweep, exactly as a reference reader walking the store sees them.)  The library
 * enforces this with stream order: update batches run on the context's stream; each waits for the last search
 * enqueued on another stream, and each search enqueued on another stream waits for the last update batch.
 * Exception: a search captured into a CUDA graph carries no such dependency -- order graph launches against
 * spl_gpu_sync yourself. */
int spl_gpu_sync(spl_gpu_ctx *ctx, spl_gpu_sync_stats *stats /* nullable */);

/* Sleep until the store looks different from the last spl_gpu_sync, or timeout_ms passes.  Replaces a
 * watcher's splinter_event_bus_open + splinter_event_bus_wait loop (splinter.c:1002-1033): the owner's
 * eventfd (H->event_bus.owner_fd / owner_pid, splinter.h:144-148) is duplicated into this process (dup in
 * the owner itself, pidfd_getfd elsewhere) and polled; H->epoch (splinter.h:186) is the gate.  Because
 * only the process that ran splinter_event_bus_init writes that fd (splinter.c:929-939: g_event_fd is
 * per process) the wait also re-reads H->epoch every few milliseconds, so writes by other processes are
 * noticed without the bus, and a store with no bus at all degrades to that timed check.
 * Returns 1 = H->epoch moved since the last sync (call spl_gpu_sync), 0 = timed out unchanged,
 * <0 = error.  Like the reference's wait it consumes the eventfd count it wakes up on.
 * splinter_unset moves neither H->epoch nor the bus (splinter.c:326-362): sweep on a timer as well. */
int spl_gpu_wait(spl_gpu_ctx *ctx, uint32_t timeout_ms);

/* dense contexts: copy rows [first_row, first_row+n_rows) from host arrays.  flags bits are
 * SPL_ROW_*; bloom/flags/epoch may be NULL (-> 0 / all candidate bits / 2). */
#define SPL_ROW_LIST 0x01u /* splinter_list candidate:      hash!=0 && val_len>0 && epoch!=0 && key[0] */
#define SPL_ROW_ENUM 0x02u /* enumerate_matches candidate:  hash!=0 && epoch!=0 && key[0]              */
#define SPL_ROW_EMB  0x04u /* a consistent embedding copy exists (else sim=0, dist=FLT_MAX)            */
int spl_gpu_dense_upload(spl_gpu_ctx *ctx, uint64_t first_row, uint64_t n_rows, const float *E,
                         const uint64_t *bloom, const uint8_t *flags, const uint64_t *epoch);

/* dense contexts: generate SURVEY 8(d)'s synthetic corpus on the device (counter RNG keyed by
 * (seed, global row, d); bit-identical to the oracle's generator). */
int spl_gpu_dense_fill_synthetic(spl_gpu_ctx *ctx, uint64_t seed_corpus, uint64_t seed_bloom);

/* read back mirror rows (tests / debugging): any output pointer may be NULL */
int spl_gpu_read_mirror(spl_gpu_ctx *ctx, uint64_t first_slot, uint64_t n, float *E,
                        uint64_t *bloom, uint8_t *flags, uint64_t *epoch);

/* Resolve what the search command prints beside the scores for one hit: the slot's key, value length
 * and type flag, as cmd_search copies them out of its snapshots (_search.c:377-383).  Store-backed
 * contexts only; reads the live store, so a slot rewritten since the last sync shows its new key. */
int spl_gpu_slot_meta(const spl_gpu_ctx *ctx, uint32_t slot_idx, char key_out[64], uint32_t *val_len,
                      uint8_t *type_flag);

/* Find the slot a key lives in: FNV-1a of the key, linear probe from hash % slots comparing the hash and then
 * the key -- the lookup every keyed call of the reference starts with (splinter.c:67-81, 535-539).  The search
 * command needs it once per query, for the scratch key it must leave out of the results (_search.c:367).
 * Store-backed contexts only; reads the live store.  0 = found (*slot_idx set), -1 / ENOENT = no such key. */
int spl_gpu_find_key(const spl_gpu_ctx *ctx, const char *key, uint32_t *slot_idx);

/* ---- search ------------------------------------------------------------------------- */

/* Host-buffer search (the reference-facing call).  q = nq x dim fp32.  out = nq x out_stride
 * records; out_stride >= k when k > 0.  counts[nq] receives the number of hits per query
 * (<= k; fewer when fewer candidates survive).  With k == 0 the call returns every survivor,
 * sorted, up to out_stride per query, and counts[] holds the TOTAL number of survivors (which
 * may exceed out_stride: call again with a larger buffer).
 * Results per query are sorted by p->order. */
int spl_gpu_search(spl_gpu_ctx *ctx, const spl_gpu_search_params *p, const float *q,
                   spl_gpu_result *out, uint32_t out_stride, uint32_t *counts);

/* How many slots survived the candidate predicate and the thresholds in each query of the most recent
 * host-buffer spl_gpu_search on this context -- the reference's "active_keys" (result_count before the
 * limit is applied, _search.c:408,415-424).  Exact for FAST and for k = 0; with SPL_MODE_EXACT and a
 * threshold flag the fused path counts with the fast scores (use k = 0 when the exact count matters). */
int spl_gpu_last_totals(const spl_gpu_ctx *ctx, uint32_t *totals, uint32_t nq);

/* Device-buffer search: d_q / d_out / d_counts are device pointers on the context's device;
 * the launches are enqueued on `stream` (a cudaStream_t; NULL = the context's stream) and the
 * call does not synchronise.  k must be in 1..SPL_GPU_KMAX_FUSED.
 *
 * Ordering: the call is ordered on the GPU after every spl_gpu_sync update batch enqueued before it and
 * before every one enqueued after it (see spl_gpu_sync), and after the previous search of this context
 * on a different stream (searches of one context share its scratch, so they run one after another
 * whatever streams they are on; use one context per concurrent search stream for overlap).
 *
 * A query the requested mode cannot answer on the device alone never comes back as "0 hits":
 * d_counts[q] = SPL_GPU_COUNT_UNANSWERED (and its records are undefined) when
 *   - SPL_MODE_BATCH_BF16 / AUTO->batched: the query's candidate list overflowed, or the bf16 scores
 *     could not prove that the re-scored candidates contain the true top-k (a crowd of near-duplicates
 *     inside the bf16 error bound);
 *   - SPL_MODE_EXACT: the selection proof failed even with the widest candidate margin.
 * Re-issue such a query with SPL_MODE_FAST (or through spl_gpu_search, which falls back by itself).
 * spl_gpu_merge_topk_device / _batch_device propagate the marker: a merged query is UNANSWERED when any of
 * its input lists is. */
#define SPL_GPU_KMAX_FUSED 256u
#define SPL_GPU_COUNT_UNANSWERED 0xFFFFFFFFu
int spl_gpu_search_device(spl_gpu_ctx *ctx, const spl_gpu_search_params *p, const float *d_q,
                          spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_counts, void *stream);

/* Merge per-shard result lists (already sorted or not) into the global top-k on the device:
 * d_lists = nlists x list_stride records, d_counts[nlists].  Used after the all-gather of the
 * per-GPU local top-k.  Enqueued on `stream`, no synchronisation. */
int spl_gpu_merge_topk_device(spl_gpu_ctx *ctx, const spl_gpu_result *d_lists, const uint32_t *d_counts,
                              uint32_t nlists, uint32_t list_stride, uint32_t order, uint32_t k,
                              spl_gpu_result *d_out, uint32_t *d_out_count, void *stream);

/* The same for a batch: d_lists = [nlists][nq][list_stride] records and d_counts = [nlists][nq], exactly what an
 * all-gather of every GPU's [nq][list_stride] result block delivers; d_out = [nq][out_stride], d_out_counts[nq].
 * One launch merges all nq queries.  (No counterpart in the reference: it has one address space.) */
int spl_gpu_merge_topk_batch_device(spl_gpu_ctx *ctx, const spl_gpu_result *d_lists, const uint32_t *d_counts,
                                    uint32_t nlists, uint32_t nq, uint32_t list_stride, uint32_t order, uint32_t k,
                                    spl_gpu_result *d_out, uint32_t out_stride, uint32_t *d_out_counts, void *stream);

/* The same merge over PACKED shard blocks, the form that needs one transfer (one all-gather, or one peer copy) per
 * shard and batch instead of two: a shard's block is its [nq][list_stride] records followed by its [nq] counts,
 * padded to a whole number of records -- SPL_GPU_PACKED_BLOCK_BYTES(nq, list_stride) bytes.  Point
 * spl_gpu_search_device's d_out at the start of the block and its d_counts at SPL_GPU_PACKED_COUNTS_OFFSET bytes
 * into it; d_packed = nlists such blocks back to back. */
#define SPL_GPU_PACKED_COUNTS_OFFSET(nq, list_stride) ((size_t)(nq) * (list_stride) * 32u)
#define SPL_GPU_PACKED_BLOCK_BYTES(nq, list_stride) (SPL_GPU_PACKED_COUNTS_OFFSET(nq, list_stride) + (((size_t)(nq) * 4u + 31u) & ~(size_t)31u))
int spl_gpu_merge_topk_packed_device(spl_gpu_ctx *ctx, const void *d_packed, uint32_t nlists, uint32_t nq,
                                     uint32_t list_stride, uint32_t order, uint32_t k, spl_gpu_result *d_out,
                                     uint32_t out_stride, uint32_t *d_out_counts, void *stream);

/* ---- the step before the search: bulk embedding of ingested text ---------------------- */

/* The reference's `ingest` command stores a document as VARTEXT chunk slots carrying label 0x200
 * (splinter_cli_cmd_ingest.c:50-51,140-172) and its splinference daemon embeds them ONE KEY AT A TIME: list the keys,
 * snapshot each, skip what is not VARTEXT or already has a vector, read the text, run the model, splinter_set_embedding,
 * check the epoch moved by exactly 2 (splinference.cpp:196-325).  This call is the bulk form of that loop:
 *   1. spl_gpu_sync, then walk the mirror's host-side metadata for the slots that are enumerate_matches candidates of
 *      label_mask (0 = every listed key), are typed VARTEXT and whose mirrored vector is still all zero;
 *   2. copy their texts out of the store's value arena under the seqlock rule (epoch even and unchanged around the copy;
 *      a busy slot is skipped and left for the next call), `batch` of them at a time;
 *   3. hand each batch to embed() -- the caller's model; its arithmetic is the caller's business (the reference's is
 *      llama.cpp + a GGUF file, un-vendored and unpinned: parity at that boundary is unpinned by construction);
 *   4. write every vector back through set_embedding() -- pass the reference's own splinter_set_embedding: the store
 *      library stays the only writer of the store -- and, like splinference, count a slot whose epoch did not move by
 *      exactly 2 as raced (its vector was written, but over newer text; the next call sees it as embedded);
 *   5. one spl_gpu_sync at the end brings all new vectors into HBM in one pass.
 * Store-backed contexts only.  Returns 0 / -1 / -2 like every other call; embed() returning non-zero aborts with -1. */
typedef int (*spl_gpu_embed_fn)(void *user, uint32_t n, const char *const *texts, const uint32_t *text_lens,
                                uint32_t dim, float *vectors_out /* n x dim */);
typedef int (*spl_gpu_set_embedding_fn)(const char *key, const float *vec);   /* = splinter_set_embedding (splinter.c:567-588) */
typedef struct spl_gpu_ingest_stats {
    uint64_t candidates;     /* slots that matched label / type / zero-vector */
    uint64_t embedded;       /* vectors written and confirmed (epoch moved by exactly 2) */
    uint64_t skipped_busy;   /* text could not be read consistently (writer active): left for the next call */
    uint64_t raced;          /* written, but the slot's epoch moved by more than 2 meanwhile */
    uint64_t write_failed;   /* set_embedding() refused (slot busy or gone) */
    uint64_t batches;
    double   gather_ms, embed_ms, write_ms, sync_ms;
} spl_gpu_ingest_stats;
int spl_gpu_ingest_embeddings(spl_gpu_ctx *ctx, uint64_t label_mask, uint32_t batch, spl_gpu_embed_fn embed, void *user,
                              spl_gpu_set_embedding_fn set_embedding, spl_gpu_ingest_stats *stats /* nullable */);

/* ---- several GPUs behind one call ---------------------------------------------------------------------------- */

/* One process, one context per GPU, each mirroring a contiguous slot range [i*ceil(N/G), ...) of the same store
 * (the reference itself is single-host, single-address-space: splinter.h:542-544).  A search runs the
 * per-GPU scans side by side, pulls each GPU's packed top-k block onto the first device with one peer copy, merges
 * there with the reference's comparator + global slot index, and returns the merged records: the same bytes a single
 * context over the whole store returns.  Queries a shard could not answer in the requested mode
 * (SPL_GPU_COUNT_UNANSWERED) are re-issued on the fp32 scan by the call itself.  devices[] = CUDA ordinals (NULL =
 * 0..n_devices-1; an ordinal may repeat: several shards on one GPU).  k in 1..SPL_GPU_KMAX_FUSED. */
typedef struct spl_gpu_group spl_gpu_group;
spl_gpu_group *spl_gpu_group_attach(const char *name_or_path, uint32_t n_devices, const int32_t *devices,
                                    const spl_gpu_opts *opts /* device, slot_begin, slot_count are set per shard */);
spl_gpu_group *spl_gpu_group_open_dense(uint64_t n_rows, uint32_t dim, uint32_t n_devices, const int32_t *devices,
                                        const spl_gpu_opts *opts);
void     spl_gpu_group_detach(spl_gpu_group *g);
uint32_t spl_gpu_group_size(const spl_gpu_group *g);
spl_gpu_ctx *spl_gpu_group_member(spl_gpu_group *g, uint32_t i);   /* shard i's context (info, counters, dense upload) */
/* all shards sweep their slot ranges at once; *total (nullable) = sums, and the slowest shard's times */
int spl_gpu_group_sync(spl_gpu_group *g, spl_gpu_sync_stats *total);
int spl_gpu_group_fill_synthetic(spl_gpu_group *g, uint64_t seed_corpus, uint64_t seed_bloom);
/* host-buffer search over all shards; arguments as spl_gpu_search */
int spl_gpu_group_search(spl_gpu_group *g, const spl_gpu_search_params *p, const float *q, spl_gpu_result *out,
                         uint32_t out_stride, uint32_t *counts);
uint64_t spl_gpu_group_reissued(const spl_gpu_group *g);   /* queries re-issued on the fp32 scan so far */

/* Allocation-free variant of the fused single-query scan.  (No counterpart in the reference; it
 * exists only because of how GPU memory is managed.)
 *
 * For callers that cannot call cudaMalloc: hosts that gave all device memory to their own arena or
 * pool, latency-critical loops that cannot take the device-wide stall an allocation implies, or
 * processes whose device memory is full by design.  The caller passes everything as device pointers
 * on the CURRENT device -- the mirror columns (E = n_rows x p->dim fp32 rows, bloom, flags with
 * SPL_ROW_* bits, epoch), the query vector(s) and the result buffers -- and the kernel's scratch
 * (per-CTA candidate lists + two counters) lives in a fixed 257 KB array inside the library image,
 * which every GPU that loads the library holds until the process exits.
 *
 *  - Limits: 1 <= p->k <= 64; p->dim in {128, 256, 512, 768, 1024}; p->mode AUTO or FAST; at most
 *    256 SMs.  Anything larger is refused with -2 / EINVAL and a message (use spl_gpu_search).
 *  - No cudaMalloc / cudaFree or other allocating call is made.  The FIRST call on a device looks
 *    the array's device address up and configures the kernel; make that first call outside any
 *    stream capture.
 *  - All calls on a device share the one array and it keeps the previous call's contents; each call
 *    clears the counters it needs at zero on `stream` before launching.  Two calls whose work
 *    overlaps on the GPU (different streams, no dependency) race on the scratch and return garbage:
 *    order them.
 *  - Work is enqueued on `stream` (a cudaStream_t, may be NULL for the default stream); the call
 *    does not synchronise.  Results: d_out[q * out_stride + i], d_counts[q], slot indices are
 *    global_row0 + row. */
int spl_gpu_scan_noalloc(const float *d_E, const uint64_t *d_bloom, const uint8_t *d_flags,
                         const uint64_t *d_epoch, uint64_t n_rows, uint64_t global_row0,
                         const spl_gpu_search_params *p, const float *d_q, spl_gpu_result *d_out,
                         uint32_t out_stride, uint32_t *d_counts, void *stream);

/* number of kernel launches this context has enqueued so far (bench.py's gpu_launches) */
uint64_t spl_gpu_launch_count(const spl_gpu_ctx *ctx);

/* last kernel's device time of the most recent spl_gpu_search in ms (CUDA events on the
 * context's stream), or < 0 if timing is off.  Enable with spl_gpu_set_timing(ctx, 1). */
int    spl_gpu_set_timing(spl_gpu_ctx *ctx, int on);
double spl_gpu_last_scan_ms(const spl_gpu_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SPLINTER_GPU_H */
