/*
 * b2o_batch.c — pthread batch drivers around the oracle (TEST INFRASTRUCTURE,
 * see b2o.h).  Used by bench.py's cpu_baseline when the reference library
 * (oracle/_ref) is unavailable, and by tests to check big batches quickly.
 * Static split: one contiguous message range per thread.
 */
#include "b2o.h"

#include <pthread.h>
#include <time.h>

#include <string.h>

enum { OP_CRC = 0, OP_COMP = 1, OP_UNCOMP = 2, OP_EQUAL = 3 };

struct job {
    int op;
    uint32_t lo, hi;
    const uint8_t* data;
    const uint64_t* off;
    const uint32_t* len;
    uint8_t* out;
    const uint64_t* out_off;
    const uint32_t* out_cap;
    uint32_t* u32_out; /* crc / out_len / ulen */
    int32_t* ok;
};

static void* worker(void* arg) {
    struct job* j = (struct job*)arg;
    uint32_t i;
    for (i = j->lo; i < j->hi; i++) {
        const uint8_t* src = j->data + j->off[i];
        if (j->op == OP_CRC) {
            j->u32_out[i] = b2o_crc32c_value(src, j->len[i]);
        } else if (j->op == OP_EQUAL) { /* message i of `data` against the same message in `out` */
            j->ok[i] = memcmp(src, j->out + j->out_off[i], j->len[i]) == 0;
        } else if (j->op == OP_COMP) {
            j->u32_out[i] = (uint32_t)b2o_snappy_compress(src, j->len[i], j->out + j->out_off[i]);
        } else {
            uint32_t ulen = 0;
            j->ok[i] = b2o_snappy_uncompress(src, j->len[i], j->out + j->out_off[i],
                                             j->out_cap[i], &ulen);
            j->u32_out[i] = ulen;
        }
    }
    return NULL;
}

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static double run(struct job proto, uint32_t n, int threads) {
    pthread_t th[512];
    struct job jobs[512];
    int t;
    double t0;
    if (threads < 1) threads = 1;
    if (threads > 512) threads = 512;
    t0 = now_s();
    for (t = 0; t < threads; t++) {
        jobs[t] = proto;
        jobs[t].lo = (uint32_t)((uint64_t)n * t / threads);
        jobs[t].hi = (uint32_t)((uint64_t)n * (t + 1) / threads);
        pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    for (t = 0; t < threads; t++) pthread_join(th[t], NULL);
    return now_s() - t0;
}

double b2o_batch_crc32c(const uint8_t* data, const uint64_t* off, const uint32_t* len,
                        uint32_t n, uint32_t* crc, int threads) {
    struct job j = {OP_CRC, 0, 0, data, off, len, 0, 0, 0, crc, 0};
    return run(j, n, threads);
}

double b2o_batch_compress(const uint8_t* data, const uint64_t* off, const uint32_t* len,
                          uint32_t n, uint8_t* out, const uint64_t* out_off,
                          uint32_t* out_len, int threads) {
    struct job j = {OP_COMP, 0, 0, data, off, len, out, out_off, 0, out_len, 0};
    return run(j, n, threads);
}

double b2o_batch_uncompress(const uint8_t* data, const uint64_t* off, const uint32_t* len,
                            uint32_t n, uint8_t* out, const uint64_t* out_off,
                            const uint32_t* out_cap, uint32_t* ulen, int32_t* ok,
                            int threads) {
    struct job j = {OP_UNCOMP, 0, 0, data, off, len, out, out_off, out_cap, ulen, ok};
    return run(j, n, threads);
}

/* Compare message i of arena a (at a_off[i]) with message i of arena b (at b_off[i]), len[i] bytes each.
 * Returns the number of messages that differ and stores the index of the first one (or n) in *first_bad.
 * The parity gates of bench.py and the full-size tests use it to check EVERY message of a batch. */
uint32_t b2o_batch_equal(const uint8_t* a, const uint64_t* a_off, const uint8_t* b, const uint64_t* b_off,
                         const uint32_t* len, uint32_t n, int32_t* scratch_ok, uint32_t* first_bad, int threads) {
    struct job j = {OP_EQUAL, 0, 0, a, a_off, len, (uint8_t*)b, b_off, 0, 0, scratch_ok};
    uint32_t i, bad = 0;
    run(j, n, threads);
    *first_bad = n;
    for (i = 0; i < n; i++)
        if (!scratch_ok[i]) {
            if (!bad) *first_bad = i;
            bad++;
        }
    return bad;
}
