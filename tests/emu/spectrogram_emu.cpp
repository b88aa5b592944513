// CPU emulation of one warp of the K1 spectrogram kernel (csrc/spectrogram_core.h): the 32 lanes run as a loop,
// shared memory is a plain array, __syncwarp() points are the loop boundaries.  Used by tests/test_spectrogram_emu.py
// to validate the index maps and butterflies against numpy, and to count the shared-memory wavefronts of every exchange
// instruction (bank-conflict check), before GPU time is spent.  Test infrastructure only.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

struct TraceRec { int kind, width; long addr; };
static std::vector<TraceRec> g_trace[32];
static int g_lane = 0;
static std::vector<int> g_extra;
static const void *g_sm_base = nullptr;
static void trace_record(int kind, const void *ptr, int width) {
    g_trace[g_lane].push_back({kind, width, (long)((const char *)ptr - (const char *)g_sm_base)});
}
#define NNTK_SPEC_TRACE(kind, ptr, width) trace_record(kind, (const void *)(ptr), width)
#include "../../nntoolkitcore_b200/csrc/spectrogram_core.h"

// wavefronts of one warp-wide shared-memory instruction: 16-byte accesses are served a quarter-warp at a time, 8-byte a
// half-warp, 4-byte the whole warp; inside a group, each distinct 128-byte-line bank (4 bytes) can serve one address.
static int wavefronts(const long *addr, int width) {
    const int group = width == 16 ? 8 : (width == 8 ? 16 : 32);
    int total = 0;
    for (int g0 = 0; g0 < 32; g0 += group) {
        std::map<int, std::vector<long>> banks;
        for (int l = g0; l < g0 + group; ++l)
            for (int b = 0; b < width; b += 4) {
                long a = addr[l] + b;
                auto &v = banks[(int)((a >> 2) & 31)];
                bool seen = false;
                for (long x : v) seen |= (x == a);
                if (!seen) v.push_back(a);
            }
        size_t worst = 1;
        for (auto &kv : banks) worst = kv.second.size() > worst ? kv.second.size() : worst;
        total += (int)worst;
    }
    return total;
}

// returns the number of EXTRA wavefronts (beyond the conflict-free minimum) over all exchange instructions
extern "C" int emu_spectrogram512_pair(const float *za, const float *zb, const float *tw /*[512][2]*/,
                                       float *pa /*[257]*/, float *pb /*[257]*/, float *zout /*[512][2] or null*/) {
    std::vector<float2> sm(NNTK_SPEC_SMEM_F2);
    g_sm_base = sm.data();
    for (auto &t : g_trace) t.clear();
    float2 v[32][16];
    SpecLaneConst c[32];
    const float2 *tw2 = reinterpret_cast<const float2 *>(tw);
    for (int l = 0; l < 32; ++l) {
        spec_lane_const(l, tw2, c[l]);
        for (int j = 0; j < 16; ++j) v[l][j] = make_float2(za[l + 32 * j], zb[l + 32 * j]);
    }
    for (int l = 0; l < 32; ++l) { spec_pass1(v[l], c[l]); }
    for (int l = 0; l < 32; ++l) { g_lane = l; spec_xchg1_store(v[l], sm.data() + spec_x1_st_base(l)); }
    for (int l = 0; l < 32; ++l) { g_lane = l; spec_xchg1_load(v[l], sm.data() + spec_x1_ld_base(l)); }
    for (int l = 0; l < 32; ++l) spec_pass2(v[l], c[l]);
    for (int l = 0; l < 32; ++l) { g_lane = l; spec_xchg2_store(v[l], sm.data() + spec_x2_st_base(l)); }
    for (int l = 0; l < 32; ++l) {
        g_lane = l;
        spec_xchg2_load(v[l], sm.data() + spec_x2_ld_base(l, 0), sm.data() + spec_x2_ld_base(l, 1));
    }
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
    // bank-conflict accounting: instruction i of every lane forms one warp-wide access
    int extra = 0;
    g_extra.clear();
    const size_t n_inst = g_trace[0].size();
    for (size_t i = 0; i < n_inst; ++i) {
        long addr[32];
        for (int l = 0; l < 32; ++l) addr[l] = g_trace[l][i].addr;
        const int w = g_trace[0][i].width;
        const int ideal = w == 16 ? 4 : (w == 8 ? 2 : 1);
        g_extra.push_back(wavefronts(addr, w) - ideal);
        extra += g_extra.back();
    }
    return extra;
}

extern "C" int emu_spectrogram512_smem_instructions(void) { return (int)g_trace[0].size(); }
extern "C" void emu_spectrogram512_conflicts(int *out) { for (size_t i = 0; i < g_extra.size(); ++i) out[i] = g_extra[i]; }
