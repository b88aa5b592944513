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
    for (int l = 0; l < 32; ++l) {
        float qa[9], qb[9];
        spec_split_regs(l, v[l], qa, qb);
        for (int slot = 0; slot < 9; ++slot) {
            if (slot == 8 && l != 0) continue;
            int k = spec_slot_bin(l, slot);
            if (k <= 256) { pa[k] = qa[slot]; pb[k] = qb[slot]; }
        }
        if (zout)
            for (int s2 = 0; s2 < 2; ++s2)
                for (int k2 = 0; k2 < 8; ++k2) {
                    int k = 64 * k2 + spec_lane_c(l, s2);
                    zout[2 * k] = v[l][8 * s2 + k2].x; zout[2 * k + 1] = v[l][8 * s2 + k2].y;
                }
    }
}
