// CPU emulation of one warp of the K1 spectrogram kernel (csrc/spectrogram_core.h): the 32 lanes run as a loop,
// shared memory is a plain array, __syncwarp() points are the loop boundaries.  Used by tests/test_spectrogram_emu.py
// to validate the index maps and butterflies against numpy before GPU time is spent.  Test infrastructure only.
#include <vector>
#include "../../nntoolkitcore_b200/csrc/spectrogram_core.h"

extern "C" void emu_spectrogram512_pair(const float *za, const float *zb, const float *tw /*[512][2]*/,
                                        float *pa /*[257]*/, float *pb /*[257]*/, float *zout /*[512][2] or null*/) {
    std::vector<float2> sm(NNTK_SPEC_SMEM_F2);
    float2 v[32][16];
    SpecLaneConst c[32];
    const float2 *tw2 = reinterpret_cast<const float2 *>(tw);
    for (int l = 0; l < 32; ++l) {
        spec_lane_const(l, tw2, c[l]);
        for (int j = 0; j < 16; ++j) v[l][j] = make_float2(za[l + 32 * j], zb[l + 32 * j]);
    }
    for (int l = 0; l < 32; ++l) { spec_pass1(v[l], c[l]); }
    for (int l = 0; l < 32; ++l) spec_xchg1_store(l, v[l], sm.data());
    for (int l = 0; l < 32; ++l) spec_xchg1_load(l, v[l], sm.data());
    for (int l = 0; l < 32; ++l) spec_pass2(v[l], c[l]);
    for (int l = 0; l < 32; ++l) spec_xchg2_store(l, v[l], sm.data());
    for (int l = 0; l < 32; ++l) spec_xchg2_load(l, v[l], sm.data());
    for (int l = 0; l < 32; ++l) spec_pass3(v[l]);
    for (int l = 0; l < 32; ++l) spec_xchg3_store(l, v[l], sm.data());
    for (int k = 0; k <= 256; ++k) spec_split_power(sm.data(), k, pa[k], pb[k]);
    if (zout)
        for (int k = 0; k < 512; ++k) { zout[2 * k] = sm[spec_pad(k)].x; zout[2 * k + 1] = sm[spec_pad(k)].y; }
}
