/*
 * b2o_corpus.c — deterministic synthetic payload corpora (TEST INFRASTRUCTURE,
 * see b2o.h).  Shapes follow the reference's own benchmark inputs:
 *   az09   : "a..z0..9" repeated   (test/brpc_snappy_compress_unittest.cpp:110-117,176-183)
 *   rand62 : uniform over [a-zA-Z0-9] (same file :203-218, there with unseeded rand())
 *   words  : RPC-body-like text from a 16-entry dictionary with sparse noise
 * Message `index` of a corpus is a pure function of (kind, index, n), so the
 * same bytes can be regenerated on any box.
 */
#include "b2o.h"

#include <pthread.h>
#include <string.h>

static const char kAlnum[] =
    "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789";

static const char* const kDict[16] = {
    "request ", "response ", "id=", "user", "name", "value", "0123", "status:ok ",
    "error ", "timestamp", "payload", "brpc", "echo", "message", "{\"k\":", "\"v\"}, "};

static inline uint32_t lcg(uint32_t s) { return s * 1664525u + 1013904223u; }

static inline uint32_t mix_seed(uint32_t seed, uint64_t index) {
    /* index 0 keeps the plain seed; others are spread by a Weyl step */
    return seed + (uint32_t)(index * 2654435761u) + (uint32_t)(index >> 32) * 40503u;
}

void b2o_corpus_fill(int kind, uint64_t index, uint8_t* out, size_t n) {
    size_t i;
    if (kind == B2O_AZ09) {
        for (i = 0; i < n; i++) {
            uint32_t r = (uint32_t)(i % 36);
            out[i] = (uint8_t)(r < 26 ? 'a' + r : '0' + (r - 26));
        }
    } else if (kind == B2O_RAND62) {
        uint32_t s = mix_seed(12345u, index);
        for (i = 0; i < n; i++) {
            s = lcg(s);
            out[i] = (uint8_t)kAlnum[(s >> 16) % 62];
        }
    } else {
        uint32_t s = mix_seed(777u, index);
        i = 0;
        while (i < n) {
            const char* w;
            s = lcg(s);
            w = kDict[(s >> 20) & 15];
            while (*w && i < n) out[i++] = (uint8_t)*w++;
            if (((s >> 8) & 7) == 0 && i < n) out[i++] = (uint8_t)('a' + (s >> 11) % 26);
        }
    }
}

struct fill_job {
    int kind;
    uint64_t first_index;
    uint32_t lo, hi;
    const uint64_t* off;
    const uint32_t* len;
    uint8_t* out;
};

static void* fill_worker(void* arg) {
    struct fill_job* j = (struct fill_job*)arg;
    uint32_t i;
    for (i = j->lo; i < j->hi; i++)
        b2o_corpus_fill(j->kind, j->first_index + i, j->out + j->off[i], j->len[i]);
    return NULL;
}

void b2o_corpus_fill_batch(int kind, uint64_t first_index, uint32_t n_msgs,
                           const uint64_t* off, const uint32_t* len,
                           uint8_t* out, int threads) {
    pthread_t th[256];
    struct fill_job jobs[256];
    int t;
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    for (t = 0; t < threads; t++) {
        jobs[t].kind = kind;
        jobs[t].first_index = first_index;
        jobs[t].lo = (uint32_t)((uint64_t)n_msgs * t / threads);
        jobs[t].hi = (uint32_t)((uint64_t)n_msgs * (t + 1) / threads);
        jobs[t].off = off;
        jobs[t].len = len;
        jobs[t].out = out;
        pthread_create(&th[t], NULL, fill_worker, &jobs[t]);
    }
    for (t = 0; t < threads; t++) pthread_join(th[t], NULL);
}
