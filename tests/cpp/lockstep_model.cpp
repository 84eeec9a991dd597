// Host build of the lock-step lane code (brpc-github-actions_b200/csrc/snappy_lockstep.cuh): the lanes of one
// group of 32 slots are run trip by trip on the CPU exactly as the kernel runs them, and the recorded copies are
// encoded by a plain serial EmitLiteral / EmitCopy (reference snappy.cc:156-232), so that
// tests/test_lockstep_model.py can check the state machine against the oracle without a GPU.
// Test infrastructure only.
#include <string.h>

#include <vector>

#include "snappy_lockstep.cuh"
#include "snappy_lockstep_dec.cuh"

using namespace b2c::ls;

static uint8_t* emit_literal(uint8_t* op, const uint8_t* src, uint32_t len) {
    const uint32_t n1 = len - 1;
    if (n1 < 60) {
        *op++ = (uint8_t)(n1 << 2);
    } else {
        uint32_t count = n1 < 256 ? 1 : n1 < 65536 ? 2 : 3;
        *op++ = (uint8_t)((59 + count) << 2);
        for (uint32_t i = 0; i < count; i++) *op++ = (uint8_t)(n1 >> (8 * i));
    }
    memcpy(op, src, len);
    return op + len;
}
static uint8_t* emit_copy_lt64(uint8_t* op, uint32_t offset, uint32_t len) {
    if (len < 12 && offset < 2048) {
        *op++ = (uint8_t)(1 + ((len - 4) << 2) + ((offset >> 8) << 5));
        *op++ = (uint8_t)(offset & 0xff);
    } else {
        *op++ = (uint8_t)(2 + ((len - 1) << 2));
        *op++ = (uint8_t)(offset & 0xff);
        *op++ = (uint8_t)(offset >> 8);
    }
    return op;
}
static uint8_t* emit_copy(uint8_t* op, uint32_t offset, uint32_t len) {
    while (len >= 68) {
        op = emit_copy_lt64(op, offset, 64);
        len -= 64;
    }
    if (len > 64) {
        op = emit_copy_lt64(op, offset, 60);
        len -= 60;
    }
    return emit_copy_lt64(op, offset, len);
}

extern "C" {

// Compress up to 32 bodies (body i = arena[off[i], off[i]+len[i]), len <= 1024) as one group of slots would.
// out: slot i at out + i * out_stride; clen[i] = fragment bytes; *trips = trips of the slowest lane;
// lane_trips[2] (nullable): lane-trips spent probing / extending, summed over lanes.
int ls_model_compress(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint32_t nb, uint8_t* out,
                      uint32_t out_stride, uint32_t* clen, uint64_t* trips, uint64_t* lane_trips) {
    if (nb > kSlots) return -1;
    std::vector<uint32_t> sm(kGroupWords, 0xdeadbeefu);  // stale shared memory must not matter
    std::vector<uint32_t> lists((size_t)kSlots * kMaxElems);
    uint32_t* tab = sm.data();
    uint32_t* body = sm.data() + kTabWords * kSlots;
    for (uint32_t i = 0; i < kTabWords * kSlots; i++) tab[i] = 0;
    FindLane L[kSlots];
    for (uint32_t l = 0; l < nb; l++) {
        if (len[l] > kBodyMax) return -2;
        for (uint32_t w = 0; w * 4 < len[l]; w++) {
            uint32_t v = 0;
            memcpy(&v, arena + off[l] + 4 * w, len[l] - 4 * w >= 4 ? 4 : len[l] - 4 * w);
            body[w * kSlots + l] = v;
        }
        L[l].init(tab + l, body + l, lists.data() + (size_t)l * kMaxElems, len[l]);
    }
    uint64_t t = 0;
    for (;;) {
        bool any = false;
        for (uint32_t l = 0; l < nb; l++)
            if (L[l].live) {
                if (lane_trips) lane_trips[L[l].ext]++;
                L[l].trip();
                any = true;
            }
        if (!any) break;
        t++;
    }
    for (uint32_t l = 0; l < nb; l++) {
        const uint8_t* src = arena + off[l];
        uint8_t* o0 = out + (size_t)l * out_stride;
        uint8_t* op = o0;
        uint32_t next_emit = 0;
        if (L[l].ne > kMaxElems) return -3;
        for (uint32_t k = 0; k < L[l].ne; k++) {
            uint32_t p, offset, ln;
            unpack_element(L[l].list[k], p, offset, ln);
            if (p > next_emit) op = emit_literal(op, src + next_emit, p - next_emit);
            op = emit_copy(op, offset, ln);
            next_emit = p + ln;
        }
        if (next_emit < len[l]) op = emit_literal(op, src + next_emit, len[l] - next_emit);
        clen[l] = (uint32_t)(op - o0);
    }
    if (trips) *trips = t;
    return 0;
}

// Decode up to 32 snappy streams (stream i = arena[off[i], off[i]+len[i]), output capacity cap[i] <= 1024) as one
// warp of the lock-step decoder would.  out: slot i at out + i * out_stride; ulen[i] = the preamble's length;
// status[i] = 0 ok / 1 corrupt / 2 too small; *trips = trips of the slowest lane.
int ls_model_decompress(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint32_t nb, const uint32_t* cap,
                        uint8_t* out, uint32_t out_stride, uint32_t* ulen, int32_t* status, uint64_t* trips) {
    using namespace b2c::lsd;
    if (nb > 32) return -1;
    std::vector<uint32_t> sm((size_t)kSlotWords * 32, 0xdeadbeefu);  // stale shared memory must not matter
    DecodeLane L[32];
    for (uint32_t l = 0; l < nb; l++) {
        if (cap[l] > kOutMax) return -2;
        L[l].init(arena + off[l], len[l], sm.data() + l, cap[l]);
    }
    uint64_t t = 0;
    for (;;) {
        bool any = false;
        for (uint32_t l = 0; l < nb; l++)
            if (L[l].live) {
                L[l].trip();
                any = true;
            }
        if (!any) break;
        t++;
    }
    for (uint32_t l = 0; l < nb; l++) {
        L[l].finish();
        ulen[l] = L[l].expected;
        status[l] = L[l].status;
        if (L[l].status == kOk)
            for (uint32_t b = 0; b < L[l].expected; b++)
                out[(size_t)l * out_stride + b] = (uint8_t)(sm[(size_t)(b >> 2) * 32 + l] >> (8 * (b & 3)));
    }
    if (trips) *trips = t;
    return 0;
}

}  // extern "C"
