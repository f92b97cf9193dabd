/* oracle/stub_embed_lib.c -- TEST INFRASTRUCTURE ONLY: the stub embedder as a batch callback (spl_gpu_embed_fn). */
#include "stub_embed.h"

int stub_embed_batch(void *user, uint32_t n, const char *const *texts, const uint32_t *lens, uint32_t dim, float *out) {
    if (user) *(uint64_t *)user += n;   /* the test counts how many texts the callback saw */
    for (uint32_t i = 0; i < n; i++) stub_embed_one(texts[i], lens[i], dim, out + (size_t)i * dim);
    return 0;
}
