// Host check of biquad_step_eft (rn_core.cuh) against the reference double formulation: for every sample
// either the hazard flag is raised or (y, m0', m1') are bit-identical.  Build + run on the CPU:
//   nvcc -O2 -fmad=false -Xcompiler -ffp-contract=off -o /tmp/eftcheck tools/ubench/biquad_eft_check.cu && /tmp/eftcheck
#include "../../formula-vad_b200/csrc/rn_core.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
using namespace fvad::rn;
static unsigned long long rs = 88172645463325252ull;
static inline unsigned long long rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
static inline float urand() { return (float)((double)(rnd() >> 11) / 9007199254740992.0 * 2.0 - 1.0); }
int main(int argc, char **argv) {
    long long N = argc > 1 ? atoll(argv[1]) : 200000000ll;
    long long hazards = 0, mism = 0, total = 0;
    for (int regime = 0; regime < 8; regime++) {
        float m0 = 0, m1 = 0, r0 = 0, r1 = 0;
        long long hz_reg = 0;
        double ph = 0;
        for (long long i = 0; i < N / 8; i++) {
            float x;
            switch (regime) {
            case 0: x = urand() * 0.03f; break;                                       // white noise
            case 1: ph += 0.013; x = 0.3f * (float)sin(ph) + urand() * 1e-3f; break;  // LF tone + noise
            case 2: x = (i / 100000) % 2 ? 0.0f : urand() * 0.5f; break;              // bursts and digital silence
            case 3: x = urand() * 1e-6f; break;                                       // very quiet
            case 4: x = urand(); break;                                               // full scale
            case 5: ph += 2.9; x = 0.9f * (float)sin(ph); break;                      // HF tone
            case 6: x = (float)((int)(rnd() % 65536) - 32768) / 32768.0f; break;      // s16 grid
            default: x = (i % 3000 < 10) ? urand() : 1e-4f * urand(); break;          // clicks
            }
            // both paths start every sample from the reference state, so each step is checked in isolation
            float a0 = r0, a1 = r1;
            bool hz = false;
            float y1 = biquad_step_eft(x, a0, a1, hz);
            float y0 = biquad_step(x, r0, r1);
            total++;
            if (hz) { hazards++; hz_reg++; continue; }
            if (memcmp(&y0, &y1, 4) || memcmp(&a0, &r0, 4) || memcmp(&a1, &r1, 4)) {
                if (mism < 10) printf("MISMATCH regime %d i %lld x %.9g: y %.9g/%.9g m0 %.9g/%.9g m1 %.9g/%.9g\n", regime, i, x, y0, y1, r0, a0, r1, a1);
                mism++;
            }
            (void)m0; (void)m1;
        }
        printf("regime %d: hazards %lld of %lld (%.4f%%)\n", regime, hz_reg, N / 8, 100.0 * hz_reg / (N / 8));
    }
    printf("total %lld samples, hazards %lld (%.4f%%), mismatches %lld\n", total, hazards, 100.0 * hazards / total, mism);
    return mism != 0;
}
