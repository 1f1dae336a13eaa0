/*
 * fvad_synth_core.h -- the seeded synthetic 48 kHz "onboard radio" generator as a pure
 * function of (seed, stream, sample index): integer hashing / phase arithmetic, one 4096-entry sine table and f32
 * operations that are individually rounded, so host and device produce the same bits (-fmad=false /
 * -ffp-contract=off).  Shared by libfvad_cuda.so (fvad_synth_fill_host / _device) and by the host-only test helper
 * oracle/libsynth_host.so, so that the CPU baseline arm of bench.py can generate its input without mapping the CUDA
 * library.  Plain C99 / C++ / CUDA.
 *
 * Content per stream: an "engine" bed (8 harmonics of a slowly wandering 80-300 Hz fundamental plus white noise,
 * about -20 dBFS), "radio speech" bursts (harmonic stack, f0 100-250 Hz, syllabic 3-5 Hz envelope, 0.3-8 s long, at
 * most one per 16 s slot), exact-zero gaps (digital silence: RNNoise's silence gate denoise.c:358-362 and
 * vol_max==0 in VAD.zig:267) and very quiet gaps.
 */
#ifndef FVAD_SYNTH_CORE_H
#define FVAD_SYNTH_CORE_H
#include <stdint.h>
#if defined(__CUDACC__)
#define FVAD_SYNTH_HD __host__ __device__ static inline
#else
#define FVAD_SYNTH_HD static inline
#endif

enum { FVAD_SYNTH_SIN_BITS = 12, FVAD_SYNTH_SIN_N = 1 << 12 };
#define FVAD_SYNTH_SLOT (16ull * 48000ull)          /* one scheduling slot = 16 s */
/* table entry i = (float)sin(2*pi*i/4096), computed in double with the host libm */
#define FVAD_SYNTH_TWO_PI (2.0 * 3.14159265358979323846)


FVAD_SYNTH_HD uint64_t fvad_synth_mix64(uint64_t z) {     // splitfvad_synth_mix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
FVAD_SYNTH_HD uint64_t fvad_synth_hash3(uint64_t seed, uint64_t a, uint64_t b) {
    return fvad_synth_mix64(fvad_synth_mix64(seed ^ (a * 0xD6E8FEB86659FD93ull)) ^ (b * 0xA0761D6478BD642Full));
}
// sine of a 32-bit phase (one turn = 2^32) by linear interpolation in the table
FVAD_SYNTH_HD float fvad_synth_sin_u32(const float *tab, uint32_t phase) {
    const uint32_t idx = phase >> (32 - FVAD_SYNTH_SIN_BITS);
    const float frac = (float)(phase & ((1u << (32 - FVAD_SYNTH_SIN_BITS)) - 1)) * (1.0f / (float)(1u << (32 - FVAD_SYNTH_SIN_BITS)));
    const float s0 = tab[idx], s1 = tab[(idx + 1) & (FVAD_SYNTH_SIN_N - 1)];
    return s0 + (s1 - s0) * frac;
}
// phase increment per sample for a frequency in milli-hertz
FVAD_SYNTH_HD uint32_t fvad_synth_inc_mhz(uint64_t mhz) { return (uint32_t)((mhz << 32) / 48000000ull); }

FVAD_SYNTH_HD float fvad_synth_sample(const float *tab, uint64_t seed, uint64_t stream, uint64_t n) {
    const uint64_t hs = fvad_synth_hash3(seed, stream, 0x5EEDull);
    // ---- gaps -------------------------------------------------------------------------------
    if ((stream & 1) && n < 24000) return 0.0f;                         // half a second of digital silence
    const uint64_t slot = n / FVAD_SYNTH_SLOT, in_slot = n - slot * FVAD_SYNTH_SLOT;
    const uint64_t hk = fvad_synth_hash3(seed, stream, 0x1000 + slot);
    const uint32_t gap_kind = (uint32_t)(hk & 7);                       // 0: exact zeros, 1: very quiet
    const uint64_t gap_len = 48000ull + (uint64_t)((hk >> 3) % 96000ull);   // 1..3 s at the slot start
    float level = 1.0f;
    if (in_slot < gap_len) {
        if (gap_kind == 0) return 0.0f;
        if (gap_kind == 1) level = 1.0e-4f;
    }
    // ---- engine bed ---------------------------------------------------------------------------
    const uint64_t f0_mhz = 80000ull + (hs % 220000ull);
    const uint32_t lfo = fvad_synth_inc_mhz(50 + ((hs >> 20) % 200)) * (uint32_t)n;      // 0.05-0.25 Hz wander
    const float wander = fvad_synth_sin_u32(tab, lfo);
    // phase modulation: up to +-12 turns of the fundamental -> a few Hz to tens of Hz of deviation
    const int64_t pm = (int64_t)(wander * (12.0f * 4294967296.0f));
    const uint32_t ph = (uint32_t)((uint64_t)fvad_synth_inc_mhz(f0_mhz) * n + (uint64_t)pm);
    float v = 0.0f;
#if defined(__CUDACC__)
#pragma unroll
#endif
    for (int h = 1; h <= 8; h++) {
        const float a = 0.003f / (float)h;
        v = v + a * fvad_synth_sin_u32(tab, ph * (uint32_t)h + (uint32_t)(hs >> (4 * h)) * 0x9E3779B9u);
    }
    const uint64_t hn = fvad_synth_hash3(seed, stream ^ 0xABCDEF, n);
    v = v + (float)(int32_t)(uint32_t)hn * (0.03f / 2147483648.0f);    // white noise, ~-35 dBFS rms
    // ---- radio speech burst ---------------------------------------------------------------------
    if (((hk >> 24) & 3) != 0) {                                         // 3 slots out of 4 carry a burst
        const uint64_t dur = 14400ull + (uint64_t)((hk >> 26) % 369600ull);       // 0.3 .. 8 s
        const uint64_t start = 4ull * 48000ull + (uint64_t)((hk >> 45) % (FVAD_SYNTH_SLOT - 4ull * 48000ull - dur));
        if (in_slot >= start && in_slot < start + dur) {
            const uint64_t t = in_slot - start;
            const uint64_t sf0 = 100000ull + ((hk >> 8) % 150000ull);
            const uint32_t sph = fvad_synth_inc_mhz(sf0) * (uint32_t)t;
            const uint32_t syl = fvad_synth_inc_mhz(3000 + ((hk >> 12) % 2000)) * (uint32_t)t;
            float env = 0.5f - 0.5f * fvad_synth_sin_u32(tab, syl + 0x40000000u);   // 0..1, starts at 0
            env = 0.15f + 0.85f * env;
            // 10 ms linear ramps at both ends
            const uint64_t edge = t < dur - t ? t : dur - t;
            if (edge < 480) env = env * ((float)edge * (1.0f / 480.0f));
            float sp = 0.0f;
#if defined(__CUDACC__)
#pragma unroll
#endif
            for (int h = 1; h <= 10; h++) {
                const uint32_t wsel = (uint32_t)(hk >> (2 * h)) & 3u;
                const float wgt = (wsel == 0 ? 0.25f : wsel == 1 ? 0.5f : wsel == 2 ? 0.75f : 1.0f) * (h <= 6 ? 0.12f : 0.05f);
                sp = sp + wgt * fvad_synth_sin_u32(tab, sph * (uint32_t)h);
            }
            v = v + env * sp;
        }
    }
    v = v * level;
    if (v > 1.0f) v = 1.0f;
    if (v < -1.0f) v = -1.0f;
    return v;
}

#endif /* FVAD_SYNTH_CORE_H */
