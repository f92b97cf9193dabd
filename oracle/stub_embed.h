/*
 * oracle/stub_embed.h -- TEST INFRASTRUCTURE ONLY.
 *
 * A deterministic stand-in for the embedding model (the reference's is llama.cpp + a Nomic GGUF, un-vendored and
 * unpinned: CMakeLists.txt:146-147): the vector is a pure function of the text bytes, unit length, dense.
 * Shared by the one-key-at-a-time loop (ref_backfill.c) and the bulk call's embed callback (libstub_embed.so), so
 * both write the same bytes for the same text.
 */
#ifndef SPL_STUB_EMBED_H
#define SPL_STUB_EMBED_H
#include <math.h>
#include <stddef.h>
#include <stdint.h>

static inline void stub_embed_one(const char *text, uint32_t len, uint32_t dim, float *out) {
    uint64_t h = 14695981039346656037ULL;
    for (uint32_t i = 0; i < len; i++) h = (h ^ (unsigned char)text[i]) * 1099511628211ULL;
    double ss = 0.0;
    for (uint32_t d = 0; d < dim; d++) {
        uint64_t z = h + 0x9E3779B97F4A7C15ULL * (d + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        z ^= z >> 31;
        out[d] = (float)((double)(z >> 11) / 9007199254740992.0 - 0.5);
        ss += (double)out[d] * out[d];
    }
    const float sc = (float)(1.0 / sqrt(ss));
    for (uint32_t d = 0; d < dim; d++) out[d] *= sc;
}
#endif
