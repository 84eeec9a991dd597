// Host build of the lock-step lane code (brpc-github-actions_b200/csrc/snappy_lockstep.cuh): the 32 lanes of one
// warp are run trip by trip on the CPU exactly as the kernels run them, so tests/test_lockstep_model.py can
// check the state machines against the oracle without a GPU.  Test infrastructure only.
#include <string.h>

#include <vector>

#include "snappy_lockstep.cuh"

using namespace b2c::ls;

extern "C" {

// Compress up to 32 bodies (body i = arena[off[i], off[i]+len[i]), len <= 1024) as one warp would.
// out: slot i at out + i * out_stride (4-byte aligned); clen[i] = fragment bytes; *trips = main-loop trips.
// mode_trips[5] (nullable): lane-trips spent in each mode, summed over lanes.
int ls_model_compress(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint32_t nb, uint8_t* out,
                      uint32_t out_stride, uint32_t* clen, uint64_t* trips, uint64_t* mode_trips) {
    if (nb > kLanes) return -1;
    std::vector<uint32_t> sm(kWarpWords, 0xdeadbeefu);  // stale shared memory must not matter
    uint32_t* tab = sm.data();
    uint32_t* body = sm.data() + kTabWords * kLanes;
    for (uint32_t i = 0; i < kTabWords * kLanes; i++) tab[i] = 0;
    Lane L[kLanes];
    for (uint32_t l = 0; l < nb; l++) {
        if (len[l] > kBodyMax) return -2;
        for (uint32_t w = 0; w * 4 < len[l]; w++) {
            uint32_t v = 0;
            memcpy(&v, arena + off[l] + 4 * w, len[l] - 4 * w >= 4 ? 4 : len[l] - 4 * w);
            body[w * kLanes + l] = v;
        }
        L[l].init(tab + l, body + l, reinterpret_cast<uint32_t*>(out + (size_t)l * out_stride), len[l]);
    }
    uint64_t t = 0;
    for (;;) {
        bool any = false;
        for (uint32_t l = 0; l < nb; l++)
            if (L[l].mode != kStop) {
                if (mode_trips) mode_trips[L[l].mode]++;
                L[l].trip();
                any = true;
            }
        if (!any) break;
        t++;
    }
    for (uint32_t l = 0; l < nb; l++) L[l].remainder_begin();
    for (;;) {
        bool any = false;
        for (uint32_t l = 0; l < nb; l++)
            if (L[l].lit_rem) {
                L[l].remainder_step();
                any = true;
            }
        if (!any) break;
    }
    for (uint32_t l = 0; l < nb; l++) clen[l] = L[l].finish();
    if (trips) *trips = t;
    return 0;
}

}  // extern "C"
