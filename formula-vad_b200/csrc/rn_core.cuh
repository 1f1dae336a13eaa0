// rn_core.cuh -- the arithmetic of RNNoise (lib/rnnoise/src) re-organised for warps.
//
// Every routine here is written ONCE as a sequence of lane phases (RN_PHASE ... RN_END).  On the
// device a phase is the body executed by the 32 lanes of one warp followed by __syncwarp(); on the
// host (tools/rn_emu, unit tests without a GPU) the same phase is a loop over lane = 0..31.  Values
// cross phases only through the warp's scratch memory, so the two executions are the same program.
//
// Exactness: the library is compiled with -fmad=false / -ffp-contract=off.  Each accumulation below
// runs in the reference's order on ONE lane (a "chain"); parallelism comes from running independent
// chains on different lanes and different frames on different warps, never from re-associating a sum.
#pragma once
#include "common.cuh"
#include <math.h>

#if defined(__CUDA_ARCH__)
#define RN_PHASE { const int lane = threadIdx.x & 31;
#define RN_END } __syncwarp();
#else
#define RN_PHASE for (int lane = 0; lane < 32; lane++) {
#define RN_END }
#endif
#define RN_HD __host__ __device__ inline

namespace fvad {
namespace rn {

// Ly = log10(1e-2 + Ex) of compute_frame_features (denoise.c:349): a float, widened to double by the double constant, the
// double log10, rounded to float.  libdevice's log10 and glibc's are both within one ulp of the true value in DOUBLE but
// neither is correctly rounded, so the two can differ in the last double bit; that changes the float only when the value
// lies within ~2^-28 of a float rounding boundary.  Parity here is EMPIRICAL: tests/test_gpu_denoiser.py sweeps this very
// function over 10^8 inputs against the host's log10 (fvad_debug_log10) and every whole-stream test goes through it.
#if defined(__CUDACC__)
__host__ __device__
#endif
inline float log10_ex(float ex) { return (float)log10(1e-2 + ex); }

constexpr int kFrame = 480, kWin = 960, kFreq = 481, kNB = 22;
constexpr int kPBuf = 1728, kPHist = kPBuf - kFrame;      // PITCH_BUF_SIZE, denoise.c:52-55
constexpr int kPMax = 768, kPMin = 60;
constexpr int kCeps = 8, kNDelta = 6, kNFeat = 42;
constexpr int kMaxCand = 15;                              // T0 plus k = 2..15 (pitch.c:460)

struct Tables {                     // pointers valid in the executing address space
    const float *half_window;       // 480   denoise.c:171-172
    const float *dct;               // 22x22 denoise.c:173-179
    const cpx *tw;                  // 960   kiss_fft.c:404-419
    const short *rev;               // 960   kiss_fft.c:314-343
    const float *frac, *omf;        // 400: j/band_size and 1 - j/band_size per bin (denoise.c:110-116)
    const float *tansig;            // 201   tansig_table.h
    double dct_scale;               // sqrt(2./22), denoise.c:193
};

__host__ __device__ constexpr int eband(int i) {          // denoise.c:73-76
    return i == 0 ? 0 : i == 1 ? 1 : i == 2 ? 2 : i == 3 ? 3 : i == 4 ? 4 : i == 5 ? 5 : i == 6 ? 6 : i == 7 ? 7 : i == 8 ? 8
         : i == 9 ? 10 : i == 10 ? 12 : i == 11 ? 14 : i == 12 ? 16 : i == 13 ? 20 : i == 14 ? 24 : i == 15 ? 28 : i == 16 ? 34
         : i == 17 ? 40 : i == 18 ? 48 : i == 19 ? 60 : i == 20 ? 78 : 100;
}

// The same table for a run-time index without branches (a switch over a lane's band index compiles to a jump table whose
// 22 targets the lanes of a warp visit one after the other).
__host__ __device__ constexpr int eband_rt(int i) {
    int e = i;
    e = i > 8 ? 2 * i - 8 : e;
    e = i > 12 ? 4 * i - 32 : e;
    e = i > 15 ? 6 * i - 62 : e;
    e = i == 18 ? 48 : e;
    e = i == 19 ? 60 : e;
    e = i == 20 ? 78 : e;
    e = i == 21 ? 100 : e;
    return e;
}
constexpr bool eband_rt_ok() {
    for (int i = 0; i < 22; i++) if (eband_rt(i) != eband(i)) return false;
    return true;
}
static_assert(eband_rt_ok(), "eband_rt must reproduce eband5ms");

// ------------------------------------------------------------------------------------------
// 960-point Opus FFT (kiss_fft.c:518-586) on a warp; data in scratch at padded slots
// ------------------------------------------------------------------------------------------
RN_HD int s960(int p) { return p + (p >> 4); }            // one pad per 16: every stage is conflict-free
constexpr int kSlots960 = 960 + 60;

RN_HD void bfly4_tw(cpx &f0, cpx &f1, cpx &f2, cpx &f3, cpx w1, cpx w2, cpx w3) {   // kiss_fft.c:139-160
    cpx s0 = c_mul(f1, w1), s1 = c_mul(f2, w2), s2 = c_mul(f3, w3);
    cpx s5 = c_sub(f0, s1);
    f0 = c_add(f0, s1);
    cpx s3 = c_add(s0, s2);
    cpx s4 = c_sub(s0, s2);
    f2 = c_sub(f0, s3);
    f0 = c_add(f0, s3);
    f1.r = s5.r + s4.i; f1.i = s5.i - s4.r;
    f3.r = s5.r - s4.i; f3.i = s5.i + s4.r;
}
RN_HD void bfly4_m1(cpx &f0, cpx &f1, cpx &f2, cpx &f3) {                           // kiss_fft.c:112-131
    cpx s0 = c_sub(f0, f2);
    f0 = c_add(f0, f2);
    cpx s1 = c_add(f1, f3);
    f2 = c_sub(f0, s1);
    f0 = c_add(f0, s1);
    s1 = c_sub(f1, f3);
    f1.r = s0.r + s1.i; f1.i = s0.i - s1.r;
    f3.r = s0.r - s1.i; f3.i = s0.i + s1.r;
}

// The input must already sit at buf[s960(rev[i])] = (1/960) * in[i]  (kiss_fft.c:579-584).
RN_HD void fft960_warp(cpx *buf, const cpx *tw) {
    RN_PHASE   // radix-4 m=1 and radix-4 m=4 (N=60, mm=16, fstride 60) on the same 16 points
        for (int b = lane; b < 60; b += 32) {
            cpx a[16];
#pragma unroll
            for (int i = 0; i < 16; i++) a[i] = buf[s960(16 * b + i)];
#pragma unroll
            for (int g = 0; g < 4; g++) bfly4_m1(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3]);
#pragma unroll
            for (int j = 0; j < 4; j++) bfly4_tw(a[j], a[j + 4], a[j + 8], a[j + 12], tw[j * 60], tw[j * 120], tw[j * 180]);
#pragma unroll
            for (int i = 0; i < 16; i++) buf[s960(16 * b + i)] = a[i];
        }
    RN_END
    RN_PHASE   // radix-4 m=16, N=15, mm=64, fstride 15
        for (int u = lane; u < 240; u += 32) {
            const int j = u & 15, base = 64 * (u >> 4) + j;
            cpx f0 = buf[s960(base)], f1 = buf[s960(base + 16)], f2 = buf[s960(base + 32)], f3 = buf[s960(base + 48)];
            bfly4_tw(f0, f1, f2, f3, tw[j * 15], tw[j * 30], tw[j * 45]);
            buf[s960(base)] = f0; buf[s960(base + 16)] = f1; buf[s960(base + 32)] = f2; buf[s960(base + 48)] = f3;
        }
    RN_END
    RN_PHASE   // radix-3 m=64, N=5, mm=192, fstride 5 (kf_bfly3, kiss_fft.c:174-226)
        const float epi3_i = tw[5 * 64].i;
        for (int u = lane; u < 320; u += 32) {
            const int k = u & 63, base = 192 * (u >> 6) + k;
            cpx f0 = buf[s960(base)], f1 = buf[s960(base + 64)], f2 = buf[s960(base + 128)];
            cpx s1 = c_mul(f1, tw[k * 5]), s2 = c_mul(f2, tw[k * 10]);
            cpx s3 = c_add(s1, s2), s0 = c_sub(s1, s2);
            f1.r = f0.r - s3.r * .5f;
            f1.i = f0.i - s3.i * .5f;
            s0.r *= epi3_i; s0.i *= epi3_i;
            f0 = c_add(f0, s3);
            f2.r = f1.r + s0.i;
            f2.i = f1.i - s0.r;
            f1.r = f1.r - s0.i;
            f1.i = f1.i + s0.r;
            buf[s960(base)] = f0; buf[s960(base + 64)] = f1; buf[s960(base + 128)] = f2;
        }
    RN_END
    RN_PHASE   // radix-5 m=192, N=1, fstride 1 (kf_bfly5, kiss_fft.c:230-305)
        const cpx ya = tw[192], yb = tw[384];
        for (int u = lane; u < 192; u += 32) {
            cpx s0 = buf[s960(u)];
            cpx s1 = c_mul(buf[s960(u + 192)], tw[u]);
            cpx s2 = c_mul(buf[s960(u + 384)], tw[2 * u]);
            cpx s3 = c_mul(buf[s960(u + 576)], tw[3 * u]);
            cpx s4 = c_mul(buf[s960(u + 768)], tw[4 * u]);
            cpx s7 = c_add(s1, s4), s10 = c_sub(s1, s4), s8 = c_add(s2, s3), s9 = c_sub(s2, s3);
            cpx o0, s5, s6, s11, s12;
            o0.r = s0.r + (s7.r + s8.r);
            o0.i = s0.i + (s7.i + s8.i);
            s5.r = s0.r + (s7.r * ya.r + s8.r * yb.r);
            s5.i = s0.i + (s7.i * ya.r + s8.i * yb.r);
            s6.r = s10.i * ya.i + s9.i * yb.i;
            s6.i = -(s10.r * ya.i + s9.r * yb.i);
            s11.r = s0.r + (s7.r * yb.r + s8.r * ya.r);
            s11.i = s0.i + (s7.i * yb.r + s8.i * ya.r);
            s12.r = s9.i * ya.i - s10.i * yb.i;
            s12.i = s10.r * yb.i - s9.r * ya.i;
            buf[s960(u)] = o0;
            buf[s960(u + 192)] = c_sub(s5, s6);
            buf[s960(u + 768)] = c_add(s5, s6);
            buf[s960(u + 384)] = c_add(s11, s12);
            buf[s960(u + 576)] = c_sub(s11, s12);
        }
    RN_END
}

// ------------------------------------------------------------------------------------------
// band chains: compute_band_energy / compute_band_corr (denoise.c:100-148).  One lane per band;
// sum[b] receives band b-1's `frac` terms first, then band b's `1-frac` terms, in bin order.
// ------------------------------------------------------------------------------------------
struct SlotSpec { const cpx *b; RN_HD cpx operator[](int k) const { return b[s960(k)]; } };
struct LinSpec { const cpx *b; RN_HD cpx operator[](int k) const { return b[k]; } };

template <class SX, class SP, bool CORR>
RN_HD float band_chain(const Tables &t, SX X, SP P, int b) {
    float s = 0;
    if (b > 0) {
        const int lo = eband_rt(b - 1) << 2, hi = eband_rt(b) << 2;
        for (int k = lo; k < hi; k++) {
            cpx x = X[k];
            float tmp;
            if (CORR) { cpx p = P[k]; tmp = x.r * p.r; tmp += x.i * p.i; }
            else { tmp = x.r * x.r; tmp += x.i * x.i; }
            s += t.frac[k] * tmp;
        }
    }
    if (b < kNB - 1) {
        const int lo = eband_rt(b) << 2, hi = eband_rt(b + 1) << 2;
        for (int k = lo; k < hi; k++) {
            cpx x = X[k];
            float tmp;
            if (CORR) { cpx p = P[k]; tmp = x.r * p.r; tmp += x.i * p.i; }
            else { tmp = x.r * x.r; tmp += x.i * x.i; }
            s += t.omf[k] * tmp;
        }
    }
    if (b == 0 || b == kNB - 1) s *= 2;
    return s;
}

// The same chain over per-bin products e[k] (= |X[k]|^2 or Re(X[k] conj P[k]), denoise.c:112-113,137-138) that all
// 32 lanes formed beforehand: only the weighted in-order accumulation stays on the band's lane.
RN_HD float band_chain_pre(const Tables &t, const float *e, int b) {
    float s = 0;
    if (b > 0) {
        const int lo = eband_rt(b - 1) << 2, hi = eband_rt(b) << 2;
        for (int k = lo; k < hi; k++) s += t.frac[k] * e[k];
    }
    if (b < kNB - 1) {
        const int lo = eband_rt(b) << 2, hi = eband_rt(b + 1) << 2;
        for (int k = lo; k < hi; k++) s += t.omf[k] * e[k];
    }
    if (b == 0 || b == kNB - 1) s *= 2;
    return s;
}
// Two such chains over the same bins (band energy of P and band correlation X.P) walked together: the weights, the
// indices and the loop are shared, each sum is still its own in-order chain.
RN_HD void band_chain_pre2(const Tables &t, const float *e0, const float *e1, int b, float &s0, float &s1) {
    float a = 0, c = 0;
    if (b > 0) {
        const int lo = eband_rt(b - 1) << 2, hi = eband_rt(b) << 2;
        for (int k = lo; k < hi; k++) { const float w = t.frac[k]; a += w * e0[k]; c += w * e1[k]; }
    }
    if (b < kNB - 1) {
        const int lo = eband_rt(b) << 2, hi = eband_rt(b + 1) << 2;
        for (int k = lo; k < hi; k++) { const float w = t.omf[k]; a += w * e0[k]; c += w * e1[k]; }
    }
    if (b == 0 || b == kNB - 1) { a *= 2; c *= 2; }
    s0 = a; s1 = c;
}
// per-bin products live in the upper half of the FFT buffer: bins 0..480 occupy slots <= 510, floats from 1040 on are free
constexpr int kBandE = 1040;

// interp_band_gain (denoise.c:150-163) for one bin < 400
RN_HD int band_of_bin(int k) {          // band index i with 4*eband[i] <= k < 4*eband[i+1], k < 400: eband's piecewise-regular
    const int e = k >> 2;               // spacing (1, 2, 4, 6, then 8/12/18/22) as selects -- no branches, no 20 compares
    int i = e;
    i = e >= 8 ? 8 + ((e - 8) >> 1) : i;
    i = e >= 16 ? 12 + ((e - 16) >> 2) : i;
    i = e >= 28 ? 15 + (e >= 34 ? 1 : 0) : i;
    i = e >= 40 ? 17 + (e >= 48 ? 1 : 0) + (e >= 60 ? 1 : 0) + (e >= 78 ? 1 : 0) : i;
    return i;
}
RN_HD float band_interp_bin(const Tables &t, const float *bandE, int k) {
    const int i = band_of_bin(k);
    return t.omf[k] * bandE[i] + t.frac[k] * bandE[i + 1];
}

RN_HD float dct_out(const Tables &t, const float *in, int i) {   // dct, denoise.c:184-195
    float sum = 0;
    for (int j = 0; j < kNB; j++) sum += in[j] * t.dct[j * kNB + i];
    return (float)(sum * t.dct_scale);
}

// window + bit-reversed scatter of a real 960-sample frame: apply_window + forward_transform's load
// (denoise.c:247-254, 212-220; kiss_fft.c:579-584).  src(i) returns sample i.
template <class SRC>
RN_HD void load_windowed(const Tables &t, cpx *buf, SRC src) {
    RN_PHASE
        const float scale = 1.f / kWin;
        float xr[kWin / 32];                                // the lane's 30 samples: every load in flight before the first use
#pragma unroll
        for (int u = 0; u < kWin / 32; u++) xr[u] = src(lane + 32 * u);
#pragma unroll
        for (int u = 0; u < kWin / 32; u++) {
            const int i = lane + 32 * u;
            const float w = t.half_window[i < kFrame ? i : kWin - 1 - i];
            float x = xr[u];
            x *= w;
            cpx v; v.r = scale * x; v.i = scale * 0.0f;
            buf[s960(t.rev[i])] = v;
        }
    RN_END
}


// ------------------------------------------------------------------------------------------
// Warp scratch layouts (floats).  The FFT buffer is reused by analysis / P-spectrum / synthesis.
// ------------------------------------------------------------------------------------------
constexpr int kScrFft = 2 * kSlots960;                     // cpx buf[1020]
constexpr int kScrAna = kScrFft + 96;                      // + misc
constexpr int kScrSyn = kScrFft + 96;                      // + misc (the spectrum itself stays in registers)
constexpr int kScrIp = 68;                                 // inner products by LIST POSITION: [0] = xx, [1 + i] = lag list[i] (<= 59 lags)
constexpr int kScrPitch = 864 + 296 + 388 + kScrIp + 64 + 64 + 8;   // xlp, xc, yyl, ip, list, misc, pad = 1752 floats: the stride is
                                                                    // 24 mod 32 banks, so the pooled chains of a CTA's four frames
                                                                    // start 8 banks apart, and four frames take 28 032 B: EIGHT CTAs
                                                                    // per SM (round 1 indexed ip by lag: 388 floats, 33 KB, six CTAs)

// frame_analysis + the energy half of compute_frame_features (denoise.c:295-308, 349-362, 363-365):
// X, Ex, E<0.04 silence flag and the cepstrum dct(Ly) with its -12/-4 offsets.
// hpb = this state's high-passed signal with kPHist samples of history in front; frame q starts at
// hpb[kPHist + q*480].
RN_HD void analysis_warp(const Tables &t, const float *hpb, int q, cpx *Xq, float *Exq, float *cepsq, int *silq,
                         float *scr) {
    cpx *buf = reinterpret_cast<cpx *>(scr);
    float *misc = scr + kScrFft;                           // [0..21] Ex, [32..53] Ly, [64..85] log10(.01 + Ex)
    const float *x0 = hpb + (kPHist - kFrame) + (size_t)q * kFrame;   // analysis_mem | in
    load_windowed(t, buf, [&](int i) { return x0[i]; });
    fft960_warp(buf, t.tw);
    float *be = scr + kBandE;
    RN_PHASE
        for (int k = lane; k < kFreq; k += 32) {
            const cpx x = buf[s960(k)];
            Xq[k] = x;
            if (k < 400) { float tmp = x.r * x.r; tmp += x.i * x.i; be[k] = tmp; }
        }
    RN_END
    RN_PHASE
        if (lane < kNB) misc[lane] = band_chain_pre(t, be, lane);
    RN_END
    RN_PHASE   // the logarithms are independent per band: 22 lanes, not one (only the follow/logMax scan is sequential)
        if (lane < kNB) { Exq[lane] = misc[lane]; misc[64 + lane] = log10_ex(misc[lane]); }
    RN_END
    RN_PHASE
        if (lane == 0) {
            float logMax = -2, follow = -2, E = 0;
            for (int i = 0; i < kNB; i++) {                // denoise.c:351-357
                float ly = misc[64 + i];
                double a = follow - 1.5;
                double inner_max = a > ly ? a : (double)ly;
                float lm7 = logMax - 7;
                ly = (float)(lm7 > inner_max ? (double)lm7 : inner_max);
                logMax = logMax > ly ? logMax : ly;
                a = follow - 1.5;
                follow = (float)(a > ly ? a : (double)ly);
                E += misc[i];
                misc[32 + i] = ly;
            }
            *silq = (E < 0.04) ? 1 : 0;                     // denoise.c:358 (float < double literal)
        }
    RN_END
    RN_PHASE
        if (lane < kNB) {
            float c = dct_out(t, misc + 32, lane);
            if (lane == 0) c -= 12;
            if (lane == 1) c -= 4;
            cepsq[lane] = c;
        }
    RN_END
}

// ------------------------------------------------------------------------------------------
// pitch analysis: pitch_downsample + pitch_search + everything of remove_doubling that does not
// depend on the previous frame (pitch.c:148-216, 283-385, 423-458; celt_lpc.c:37-88, 198-279).
// Output: the candidate table the sequential decision (pitch_decide) walks.
// ------------------------------------------------------------------------------------------
struct PitchCand {                 // per frame
    int n;                         // candidates: index 0 is T0, index c is k = c+1
    short T[kMaxCand];
    signed char off[kMaxCand];     // pseudo-interpolation offset if this candidate wins (pitch.c:511-517)
    float g[kMaxCand], xy[kMaxCand], yy[kMaxCand];
};

// find_best_pitch, pitch.c:46-104.  Syy0 = 1 + sum_{j<len} y(j)^2, accumulated in that order by the caller.
template <class YF>
RN_HD void best_pitch2(const float *xcorr, YF y, int len, int max_pitch, float Syy0, int *best) {
    float Syy = Syy0;
    float best_num0 = -1, best_num1 = -1, best_den0 = 0, best_den1 = 0;
    best[0] = 0; best[1] = 1;
    for (int i = 0; i < max_pitch; i++) {
        if (xcorr[i] > 0) {
            float x16 = xcorr[i];
            x16 *= 1e-12f;
            const float num = x16 * x16;
            if (num * best_den1 > best_num1 * Syy) {
                if (num * best_den0 > best_num0 * Syy) {
                    best_num1 = best_num0; best_den1 = best_den0; best[1] = best[0];
                    best_num0 = num; best_den0 = Syy; best[0] = i;
                } else {
                    best_num1 = num; best_den1 = Syy; best[1] = i;
                }
            }
        }
        Syy += y(i + len) * y(i + len) - y(i) * y(i);
        Syy = Syy < 1 ? 1 : Syy;
    }
}

RN_HD void ld4(const float *p, float *v) {                 // four consecutive floats, p 16-byte aligned
#if defined(__CUDA_ARCH__)
    const float4 q = *reinterpret_cast<const float4 *>(p);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
#else
    v[0] = p[0]; v[1] = p[1]; v[2] = p[2]; v[3] = p[3];
#endif
}

RN_HD float pitch_gain(float xy, float xx, float yy) {     // compute_pitch_gain, pitch.c:416-419
    return (float)(xy / sqrt((double)(1 + xx * yy)));
}

// A block of NT = 32*F threads analyses F frames together.  Work that fills a warp (down-sampling, the FIR, the
// coarse correlation) stays "warp f owns frame f"; the chains that occupy only a few lanes per frame (5 autocorrelation
// lags, the best-pitch scans, 10 fine lags, the energy recurrences, the decision prep) are gathered from all F frames
// onto as few warps as possible, and independent chains share a phase -- a warp costs issue slots for its longest lane
// however few lanes are active.  Every chain is still one lane running the reference's order.  F = 1 on the host.
#if defined(__CUDA_ARCH__)
#define RN_BLK_PHASE(NT) { const int tid = threadIdx.x;
#define RN_BLK_END } __syncthreads();
#else
#define RN_BLK_PHASE(NT) for (int tid = 0; tid < (NT); tid++) {
#define RN_BLK_END }
#endif

template <int F>
RN_HD void pitch_blk(const float *const *xs, PitchCand *const *outs, float *scr0) {
    constexpr int NT = 32 * F;
    // per frame f (scr = scr0 + f*kScrPitch): xlp 864 | xc 296 | yyl 388 | ip 68 | list 64 ints | misc 16 | imisc 48 ints
    // misc: [0..4] ac, [8..12] num, [13] Syy of the coarse scan, [14] Syy of the fine scan
    // imisc: [0] best0 [1] best1 [2] T0 [3] nlist [4] ncand; [5..] candT, candTb
#define RN_SCR(f) (scr0 + (size_t)(f) * kScrPitch)
#define RN_XC(f) (RN_SCR(f) + 864)
#define RN_YYL(f) (RN_SCR(f) + 864 + 296)
#define RN_IP(f) (RN_SCR(f) + 864 + 296 + 388)
#define RN_LIST(f) (reinterpret_cast<int *>(RN_SCR(f) + 864 + 296 + 388 + kScrIp))
#define RN_MISC(f) (RN_SCR(f) + 864 + 296 + 388 + kScrIp + 64)
#define RN_IMISC(f) (reinterpret_cast<int *>(RN_SCR(f) + 864 + 296 + 388 + kScrIp + 64 + 16))
    RN_BLK_PHASE(NT)   // pitch_downsample, pitch.c:171-173 (warp f owns frame f)
        const int f = tid >> 5, lane = tid & 31;
        const float *x = xs[f];                            // pitch_buf of the frame (1728 samples)
        float *xlp = RN_SCR(f);
        for (int i = lane; i < 864; i += 32)
            xlp[i] = (i == 0) ? .5f * (.5f * x[1] + x[0]) : .5f * (.5f * (x[2 * i - 1] + x[2 * i + 1]) + x[2 * i]);
    RN_BLK_END
    RN_BLK_PHASE(NT)   // _celt_autocorr lag 0..4: fastN = 860 in-order chain, then the tail (celt_lpc.c:261-267)
        if (tid < 5 * F) {
            const int f = tid / 5, k = tid - 5 * f;
            const float *xlp = RN_SCR(f);
            float a = 0;
            for (int j = 0; j < 860; j += 4) {              // the lag-0 operand as one 16-byte load per four steps
                float xv[4];
                ld4(xlp + j, xv);
#pragma unroll
                for (int u = 0; u < 4; u++) a = a + xv[u] * xlp[j + u + k];
            }
            float d = 0;
            for (int i = k + 860; i < 864; i++) d = d + xlp[i] * xlp[i - k];
            RN_MISC(f)[k] = a + d;
        }
    RN_BLK_END
    RN_BLK_PHASE(NT)
        if (tid < F) {
            float *misc = RN_MISC(tid);
            float ac[5];
            for (int i = 0; i < 5; i++) ac[i] = misc[i];
            ac[0] *= 1.0001f;                              // pitch.c:187
            for (int i = 1; i <= 4; i++) ac[i] -= ac[i] * (.008f * i) * (.008f * i);   // pitch.c:196
            float lpc[4] = {0, 0, 0, 0};
            float error = ac[0];
            if (ac[0] != 0) {                              // _celt_lpc, celt_lpc.c:55-82
                for (int i = 0; i < 4; i++) {
                    float rr = 0;
                    for (int j = 0; j < i; j++) rr += lpc[j] * ac[i - j];
                    rr += ac[i + 1];
                    const float r = -rr / error;
                    lpc[i] = r;
                    for (int j = 0; j < (i + 1) >> 1; j++) {
                        const float t1 = lpc[j], t2 = lpc[i - 1 - j];
                        lpc[j] = t1 + r * t2;
                        lpc[i - 1 - j] = t2 + r * t1;
                    }
                    error = error - (r * r) * error;
                    if (error < .001f * ac[0]) break;
                }
            }
            float tmp = 1.0f;
            for (int i = 0; i < 4; i++) { tmp = .9f * tmp; lpc[i] = lpc[i] * tmp; }   // pitch.c:201-205
            const float c1 = .8f;
            misc[8] = lpc[0] + .8f;
            misc[9] = lpc[1] + c1 * lpc[0];
            misc[10] = lpc[2] + c1 * lpc[1];
            misc[11] = lpc[3] + c1 * lpc[2];
            misc[12] = c1 * lpc[3];
        }
    RN_BLK_END
    // celt_fir5 (pitch.c:106-145) in place: taps read the ORIGINAL x_lp, so compute all, then store.
#if defined(__CUDA_ARCH__)
    float yv[27];
#define RN_YV(r) yv[r]
#else
    static thread_local float yv_all[NT][27];
#define RN_YV(r) yv_all[tid][r]
#endif
    RN_BLK_PHASE(NT)
        const int f = tid >> 5, lane = tid & 31;
        const float *xlp = RN_SCR(f), *misc = RN_MISC(f);
        const float n0 = misc[8], n1 = misc[9], n2 = misc[10], n3 = misc[11], n4 = misc[12];
#pragma unroll
        for (int r = 0; r < 27; r++) {
            const int i = lane + 32 * r;
            float sum = xlp[i];
            sum = sum + n0 * (i >= 1 ? xlp[i - 1] : 0.0f);
            sum = sum + n1 * (i >= 2 ? xlp[i - 2] : 0.0f);
            sum = sum + n2 * (i >= 3 ? xlp[i - 3] : 0.0f);
            sum = sum + n3 * (i >= 4 ? xlp[i - 4] : 0.0f);
            sum = sum + n4 * (i >= 5 ? xlp[i - 5] : 0.0f);
            RN_YV(r) = sum;
        }
    RN_BLK_END
    // ... and the 4x-decimated signal of the coarse search (x_lp4 = y_lp4 = every second sample, pitch.c:322-325) goes to a
    // contiguous copy x4[0..431] in the still unused yyl|ip area: the coarse correlation then reads consecutive floats
    // (stride-2 reads of xlp were 2-way bank conflicts on every load).
    RN_BLK_PHASE(NT)
        const int f = tid >> 5, lane = tid & 31;
        float *xlp = RN_SCR(f), *x4 = RN_YYL(f);
#pragma unroll
        for (int r = 0; r < 27; r++) {
            xlp[lane + 32 * r] = RN_YV(r);
            if (!(lane & 1)) x4[(lane >> 1) + 16 * r] = RN_YV(r);
        }
    RN_BLK_END
#undef RN_YV
    RN_BLK_PHASE(NT)   // coarse search, 4x decimation: celt_pitch_xcorr(x_lp4, y_lp4, xcorr, 240, 147) (pitch.c:335)
        const int f = tid >> 5, lane = tid & 31;
        const float *x4 = RN_YYL(f);
        float *xc = RN_XC(f);
        // Lane l < 30 owns the five CONSECUTIVE lags 5l..5l+4 (147 needed) and walks j once for all five, each chain still
        // accumulated over j in order: the five y values of step j+1 are four of step j's plus one new load (a register
        // window), x arrives as one broadcast 16-byte load per four steps.  Lane 31 runs the coarse scan's
        // Syy = 1 + sum y(j)^2 (pitch.c:57-58) in the same instruction stream: its window starts at lag 0 and its "x" is
        // the window's own first element, so s[0] += y(j)*y(j).
        const int base = lane == 31 ? 0 : 5 * lane;
        float s[5] = {lane == 31 ? 1.0f : 0.0f, 0, 0, 0, 0}, w[5];
#pragma unroll
        for (int r = 0; r < 5; r++) w[r] = x4[base + r];
        for (int j0 = 0; j0 < 240; j0 += 20) {
            float xv[20];
#pragma unroll
            for (int q = 0; q < 5; q++) ld4(x4 + 192 + j0 + 4 * q, xv + 4 * q);
#pragma unroll
            for (int u = 0; u < 20; u++) {
                const float xl = lane == 31 ? w[0] : xv[u];
#pragma unroll
                for (int r = 0; r < 5; r++) s[r] = s[r] + xl * w[r];
#pragma unroll
                for (int r = 0; r < 4; r++) w[r] = w[r + 1];
                w[4] = x4[base + j0 + u + 5];                // <= 154 + 244 < 432: in bounds for every lane
            }
        }
#pragma unroll
        for (int r = 0; r < 5; r++)
            if (lane < 30 && 5 * lane + r < 147) xc[5 * lane + r] = s[r];
        if (lane == 31) RN_MISC(f)[13] = s[0];
    RN_BLK_END
    RN_BLK_PHASE(NT)
        if (tid < F) {
            const float *xlp = RN_SCR(tid);
            int best[2];
            best_pitch2(RN_XC(tid), [&](int j) { return xlp[2 * j]; }, 240, 147, RN_MISC(tid)[13], best);
            RN_IMISC(tid)[0] = best[0]; RN_IMISC(tid)[1] = best[1];
        }
    RN_BLK_END
    RN_BLK_PHASE(NT)
        const int f = tid >> 5, lane = tid & 31;
        float *xc = RN_XC(f);
        for (int i = lane; i < 294; i += 32) xc[i] = 0;    // pitch.c:352
    RN_BLK_END
    RN_BLK_PHASE(NT)   // fine search, 2x decimation (pitch.c:349-366): only lags within 2 of a coarse best
        if (tid < 10 * F) {
            const int f = tid / 10, l = tid - 10 * f;
            const float *xlp = RN_SCR(f);
            const int b0 = RN_IMISC(f)[0], b1 = RN_IMISC(f)[1];
            const int i = l < 5 ? 2 * b0 + l - 2 : 2 * b1 + l - 7;
            if (i >= 0 && i < 294) {
                float s = 0;
                for (int j = 0; j < 480; j += 4) {
                    float xv[4];
                    ld4(xlp + 384 + j, xv);
#pragma unroll
                    for (int u = 0; u < 4; u++) s = s + xv[u] * xlp[i + j + u];
                }
                RN_XC(f)[i] = s < -1 ? -1 : s;
            }
        } else if (tid < 11 * F) {                         // the fine scan's Syy (pitch.c:57-58), same length of chain
            const int f = tid - 10 * F;
            const float *xlp = RN_SCR(f);
            float Syy = 1;
            for (int j = 0; j < 480; j++) Syy = Syy + xlp[j] * xlp[j];
            RN_MISC(f)[14] = Syy;
        } else if (tid < 12 * F) {                         // dual_inner_prod's xx, pitch.c:446
            const int f = tid - 11 * F;
            const float *xlp = RN_SCR(f);
            float xx = 0;
            for (int i = 0; i < 480; i++) xx = xx + xlp[384 + i] * xlp[384 + i];
            RN_IP(f)[0] = xx;
            RN_YYL(f)[0] = xx;
        }
    RN_BLK_END
    RN_BLK_PHASE(NT)
        if (tid < F) {
            const int f = tid;
            const float *xlp = RN_SCR(f), *xc = RN_XC(f);
            int *imisc = RN_IMISC(f), *list = RN_LIST(f);
            int best[2] = {imisc[0], imisc[1]};
            best_pitch2(xc, [&](int j) { return xlp[j]; }, 480, 294, RN_MISC(f)[14], best);
            int offset = 0;
            if (best[0] > 0 && best[0] < 294 - 1) {        // pitch.c:371-384
                const float a = xc[best[0] - 1], b = xc[best[0]], c = xc[best[0] + 1];
                if ((c - a) > .7f * (b - a)) offset = 1;
                else if ((a - c) > .7f * (b - c)) offset = -1;
            }
            int pitch_index = kPMax - (2 * best[0] - offset);   // denoise.c:330
            // remove_doubling's frame-local part (pitch.c:436-445, 460-478)
            int T0 = pitch_index / 2;
            if (T0 >= 384) T0 = 383;
            imisc[2] = T0;
            unsigned need[13];
            for (int w = 0; w < 13; w++) need[w] = 0;
            auto mark = [&](int lag) { need[lag >> 5] |= 1u << (lag & 31); };
            mark(T0 - 1); mark(T0); mark(T0 + 1);
            int nc = 1;
            const int second_check[16] = {0, 0, 3, 2, 3, 2, 5, 2, 3, 2, 3, 2, 5, 2, 3, 2};
            for (int k = 2; k <= 15; k++) {
                const int T1 = (2 * T0 + k) / (2 * k);
                if (T1 < 30) break;
                int T1b;
                if (k == 2) T1b = (T1 + T0 > 384) ? T0 : T0 + T1;
                else T1b = (2 * second_check[k] * T0 + k) / (2 * k);
                imisc[5 + 2 * nc] = T1; imisc[6 + 2 * nc] = T1b;
                mark(T1 - 1); mark(T1); mark(T1 + 1); mark(T1b);
                nc++;
            }
            imisc[4] = nc;
            int nl = 0;
            for (int w = 0; w < 13; w++)                   // ascending lags
                for (unsigned m = need[w]; m; m &= m - 1) {
                    int bit = 0;
                    for (unsigned lowest = m & (0u - m); lowest > 1; lowest >>= 1) bit++;
                    const int lag = 32 * w + bit;
                    if (lag >= 1 && lag <= 384) list[nl++] = lag;
                }
            imisc[3] = nl;
        } else if (tid >= NT - F) {                        // pitch.c:449-453, on the last warp while warp 0 scans
            const int f = NT - 1 - tid;
            const float *xlp = RN_SCR(f);
            float *yyl = RN_YYL(f);
            float yy = yyl[0];
            for (int i = 1; i <= 384; i++) {
                yy = yy + xlp[384 - i] * xlp[384 - i] - xlp[384 + 480 - i] * xlp[384 + 480 - i];
                yyl[i] = yy < 0 ? 0 : yy;
            }
        }
    RN_BLK_END
    RN_BLK_PHASE(NT)   // inner products by lag: ip[lag] = sum_i x[i]*x[i-lag], x = x_lp + 384, N = 480; all frames' lags pooled
        int total = 0;
#pragma unroll
        for (int f = 0; f < F; f++) total += RN_IMISC(f)[3];
        for (int item = tid; item < total; item += NT) {
            int f = 0, idx = item;
#pragma unroll
            for (int g = 0; g < F - 1; g++) {
                const int nl = RN_IMISC(g)[3];
                if (f == g && idx >= nl) { idx -= nl; f = g + 1; }
            }
            const float *xlp = RN_SCR(f);
            const int lag = RN_LIST(f)[idx];
            float s = 0;
            for (int i = 0; i < 480; i += 4) {
                float xv[4];
                ld4(xlp + 384 + i, xv);
#pragma unroll
                for (int u = 0; u < 4; u++) s = s + xv[u] * xlp[384 + i + u - lag];
            }
            RN_IP(f)[1 + idx] = s;                         // by list position; ip_at() finds a lag's position
        }
    RN_BLK_END
    RN_BLK_PHASE(NT)
        if (tid < 16 * F) {
            const int f = tid >> 4, c = tid & 15;
            const int *imisc = RN_IMISC(f);
            const float *ipl = RN_IP(f), *yyl = RN_YYL(f);
            const int *list = RN_LIST(f);
            const int nl = imisc[3];
            // the inner product of a lag that the list holds (every lag asked for below was marked; ascending list)
            auto ip_at = [&](int lag) {
                int lo = 0, hi = nl - 1;
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (list[mid] < lag) lo = mid + 1; else hi = mid; }
                return ipl[1 + lo];
            };
            PitchCand *out = outs[f];
            const int nc = imisc[4];
            if (out != nullptr && c < nc) {
                const int T0 = imisc[2];
                const float xx = ipl[0];
                int T; float xy, yy;
                if (c == 0) { T = T0; xy = ip_at(T0); yy = yyl[T0]; }
                else {
                    const int T1 = imisc[5 + 2 * c], T1b = imisc[6 + 2 * c];
                    T = T1;
                    xy = .5f * (ip_at(T1) + ip_at(T1b));       // pitch.c:479-480
                    yy = .5f * (yyl[T1] + yyl[T1b]);
                }
                out->T[c] = (short)T;
                out->xy[c] = xy; out->yy[c] = yy;
                out->g[c] = pitch_gain(xy, xx, yy);
                const float x0 = ip_at(T - 1), x1 = ip_at(T), x2 = ip_at(T + 1);       // pitch.c:509-517
                int off;
                if ((x2 - x0) > .7f * (x1 - x0)) off = 1;
                else if ((x0 - x2) > .7f * (x1 - x2)) off = -1;
                else off = 0;
                out->off[c] = (signed char)off;
            }
            if (out != nullptr && c == 0) out->n = nc;
        }
    RN_BLK_END
#undef RN_SCR
#undef RN_XC
#undef RN_YYL
#undef RN_IP
#undef RN_LIST
#undef RN_MISC
#undef RN_IMISC
}

// The sequential part of remove_doubling (pitch.c:459-525): walks the candidate table with the
// previous frame's period and gain.  Returns the pitch index; *gain_out is the new last_gain.
RN_HD int pitch_decide(const PitchCand &c, int prev_period, float prev_gain, float *gain_out) {
    const int T0 = c.T[0];
    const float g0 = c.g[0];
    int T = T0, off = c.off[0];
    float g = g0, best_xy = c.xy[0], best_yy = c.yy[0];
    prev_period /= 2;
    for (int i = 1; i < c.n; i++) {
        const int k = i + 1, T1 = c.T[i];
        const float g1 = c.g[i];
        float cont;
        const int d = T1 - prev_period, ad = d < 0 ? -d : d;
        if (ad <= 1) cont = prev_gain;
        else if (ad <= 2 && 5 * k * k < T0) cont = .5f * prev_gain;
        else cont = 0;
        float a = .7f * g0 - cont;
        float thresh = .3f > a ? .3f : a;
        if (T1 < 90) { a = .85f * g0 - cont; thresh = .4f > a ? .4f : a; }
        if (g1 > thresh) { best_xy = c.xy[i]; best_yy = c.yy[i]; T = T1; g = g1; off = c.off[i]; }
    }
    best_xy = best_xy < 0 ? 0 : best_xy;
    float pg;
    if (best_yy <= best_xy) pg = 1.0f;
    else pg = best_xy / (best_yy + 1);
    if (pg > g) pg = g;
    int idx = 2 * T + off;
    if (idx < kPMin) idx = kPMin;
    *gain_out = pg;
    return idx;
}


// ------------------------------------------------------------------------------------------
// pitch-lagged spectrum and the pitch-correlation features (denoise.c:337-348)
// featB[0..5] = features[34..39], featB[6] = features[40]
// ------------------------------------------------------------------------------------------
RN_HD void pfeat_warp(const Tables &t, const float *hpb, int q, int pitch_index, const cpx *Xq, const float *Exq,
                      cpx *Pq, float *Epq, float *Expq, float *featB, float *scr) {
    cpx *buf = reinterpret_cast<cpx *>(scr);
    float *misc = scr + kScrFft;                           // [0..21] Ep, [32..53] Exp
    const float *p0 = hpb + (size_t)q * kFrame + (kPBuf - kWin - pitch_index);   // denoise.c:337-338
    load_windowed(t, buf, [&](int i) { return p0[i]; });
    fft960_warp(buf, t.tw);
    float *be = scr + kBandE, *bc = be + 400;
    RN_PHASE
        cpx xr[13];                                         // bins < 400 of X (from global memory), all loads in flight at once
#pragma unroll
        for (int u = 0; u < 13; u++) { const int k = lane + 32 * u; if (k < 400) xr[u] = Xq[k]; }
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int k = lane + 32 * u;
            if (k < kFreq) {
                const cpx pv = buf[s960(k)];
                Pq[k] = pv;
                if (u < 13 && k < 400) {
                    const cpx x = xr[u < 13 ? u : 0];
                    float tmp = pv.r * pv.r; tmp += pv.i * pv.i; be[k] = tmp;
                    tmp = x.r * pv.r; tmp += x.i * pv.i; bc[k] = tmp;
                }
            }
        }
    RN_END
    RN_PHASE
        if (lane < kNB) {
            float ep, exp_;
            band_chain_pre2(t, be, bc, lane, ep, exp_);
            exp_ = (float)(exp_ / sqrt(.001 + Exq[lane] * ep));        // denoise.c:343
            misc[lane] = ep; misc[32 + lane] = exp_;
            Epq[lane] = ep; Expq[lane] = exp_;
        }
    RN_END
    RN_PHASE
        if (lane < kNDelta) {
            float v = dct_out(t, misc + 32, lane);                     // denoise.c:344-347
            if (lane == 0) v = (float)(v - 1.3);
            if (lane == 1) v = (float)(v - 0.9);
            featB[lane] = v;
        }
        if (lane == 6) featB[6] = (float)(.01 * (pitch_index - 300));   // denoise.c:348
    RN_END
}

// ------------------------------------------------------------------------------------------
// cepstral ring features (denoise.c:366-398) for the non-silent frame whose global rank among
// non-silent frames is `rank`; cc points at its own entry of the compacted cepstrum list, so that
// cc - k*22 is the cepstrum k non-silent frames earlier (zeros before the first).
// ------------------------------------------------------------------------------------------
RN_HD void cepsfeat_warp(const float *cc, long long rank, const float *featB, float *feat, float *scr) {
    float *dist = scr;            // 64
    float *mind = scr + 64;       // 8
    const int memid = (int)(rank & 7);
    RN_PHASE
        for (int p = lane; p < 64; p += 32) {
            const int i = p >> 3, j = p & 7;
            const float *ci = cc - ((memid - i + 8) & 7) * kNB;        // slot i holds rank - ((memid-i) mod 8)
            const float *cj = cc - ((memid - j + 8) & 7) * kNB;
            float d = 0;
            for (int k = 0; k < kNB; k++) {
                const float tmp = ci[k] - cj[k];
                d += tmp * tmp;
            }
            dist[p] = d;
        }
    RN_END
    RN_PHASE
        if (lane < kCeps) {
            float m = 1e15f;
            for (int j = 0; j < kCeps; j++)
                if (j != lane) m = m < dist[lane * 8 + j] ? m : dist[lane * 8 + j];
            mind[lane] = m;
        }
        const float *c0 = cc, *c1 = cc - kNB, *c2 = cc - 2 * kNB;
        if (lane < kNB) {
            if (lane < kNDelta) {
                feat[lane] = c0[lane] + c1[lane] + c2[lane];
                feat[kNB + lane] = c0[lane] - c2[lane];
                feat[kNB + kNDelta + lane] = c0[lane] - 2 * c1[lane] + c2[lane];
            } else {
                feat[lane] = c0[lane];
            }
        }
        if (lane < 7) feat[kNB + 2 * kNDelta + lane] = featB[lane];
    RN_END
    RN_PHASE
        if (lane == 0) {
            float sv = 0;
            for (int i = 0; i < kCeps; i++) sv += mind[i];
            feat[kNB + 3 * kNDelta + 1] = (float)(sv / kCeps - 2.1);
        }
    RN_END
}

// ------------------------------------------------------------------------------------------
// pitch_filter + gains + frame_synthesis (denoise.c:422-456, 475-490, 401-408, 227-245).
// g_rnn = the RNN's band gains, g_sm = after the 0.6*lastg floor (computed by the gain scan).
// head = first 480 windowed samples, tail = last 480 (the next frame's synthesis_mem).
// ------------------------------------------------------------------------------------------
RN_HD void synth_warp(const Tables &t, const cpx *Xq, const cpx *Pq, const float *Exq, const float *Epq,
                      const float *Expq, const float *g_rnn, const float *g_sm, int silence, float *head,
                      float *tail, float *scr) {
    // Round 2: the spectrum lives in REGISTERS between the phases (lane l owns bins l + 32u, 16 per lane) instead of a
    // linear copy X[481] in shared memory: the pitch filter, the band normalisation and the gains are per-bin products, only
    // the band energies need the bins of others (|X|^2 staged in the idle FFT buffer, as before), and the Hermitian
    // extension + bit-reversed load of the inverse transform is written straight from the registers (bin k and its mirror
    // 960-k).  962 floats of scratch less per warp: 8.5 KB instead of 12.4 KB, six CTAs per SM instead of four.
    cpx *buf = reinterpret_cast<cpx *>(scr);
    float *misc = scr + kScrFft;                           // [0..21] r, [32..53] norm, [64..85] g_sm copy
#if defined(__CUDA_ARCH__)
    cpx xr[16];
#define RN_XR(u) xr[u]
#else
    static thread_local cpx xr_all[32][16];
#define RN_XR(u) xr_all[lane][u]
#endif
    RN_PHASE
        {
#pragma unroll
            for (int u = 0; u < 16; u++) { const int k = lane + 32 * u; if (k < kFreq) RN_XR(u) = Xq[k]; }
        }
        if (!silence && lane < kNB) {
            const float e = Expq[lane], g = g_rnn[lane];
            float r;
            if (e > g) r = 1;
            else r = (float)((e * e) * (1 - (g * g)) / (.001 + (g * g) * (1 - (e * e))));
            float c = 0 > r ? 0 : r;
            c = 1 < c ? 1 : c;
            r = (float)sqrt((double)c);
            r = (float)(r * sqrt(Exq[lane] / (1e-8 + Epq[lane])));
            misc[lane] = r;
            misc[64 + lane] = g_sm[lane];
        }
    RN_END
    if (!silence) {
        RN_PHASE   // pitch filter (denoise.c:422-456) and the per-bin |X|^2 of the filtered spectrum
            cpx pr[16];            // 16 loads in flight before the first use
#pragma unroll
            for (int u = 0; u < 16; u++) { const int k = lane + 32 * u; if (k < kFreq) pr[u] = Pq[k]; }
#pragma unroll
            for (int u = 0; u < 16; u++) {
                const int k = lane + 32 * u;
                if (k < kFreq) {
                    const float rf = k < 400 ? band_interp_bin(t, misc, k) : 0.0f;
                    cpx x = RN_XR(u);
                    x.r += rf * pr[u].r;
                    x.i += rf * pr[u].i;
                    RN_XR(u) = x;
                    if (k < 400) { float tmp = x.r * x.r; tmp += x.i * x.i; scr[k] = tmp; }
                }
            }
        RN_END
        RN_PHASE
            if (lane < kNB) {
                const float ne = band_chain_pre(t, scr, lane);
                misc[32 + lane] = (float)sqrt(Exq[lane] / (1e-8 + ne));
            }
        RN_END
        RN_PHASE
#pragma unroll
            for (int u = 0; u < 16; u++) {
                const int k = lane + 32 * u;
                if (k < kFreq) {
                    const float nf = k < 400 ? band_interp_bin(t, misc + 32, k) : 0.0f;
                    const float gf = k < 400 ? band_interp_bin(t, misc + 64, k) : 0.0f;
                    cpx x = RN_XR(u);
                    x.r *= nf; x.i *= nf;
                    x.r *= gf; x.i *= gf;
                    RN_XR(u) = x;
                }
            }
        RN_END
    }
    RN_PHASE   // inverse_transform's Hermitian extension + the forward FFT's scaled bit-reversed load
        const float scale = 1.f / kWin;
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int k = lane + 32 * u;
            if (k < kFreq) {
                const cpx x = RN_XR(u);
                cpx o; o.r = scale * x.r; o.i = scale * x.i;
                buf[s960(t.rev[k])] = o;
                if (k > 0 && k < kFreq - 1) {              // the mirror bin 960 - k takes the conjugate
                    cpx m; m.r = scale * x.r; m.i = scale * -x.i;
                    buf[s960(t.rev[kWin - k])] = m;
                }
            }
        }
    RN_END
#undef RN_XR
    fft960_warp(buf, t.tw);
    RN_PHASE
        for (int i = lane; i < kWin; i += 32) {
            float v = kWin * buf[s960(i == 0 ? 0 : kWin - i)].r;        // denoise.c:240-244
            v *= t.half_window[i < kFrame ? i : kWin - 1 - i];
            if (i < kFrame) head[i] = v; else tail[i - kFrame] = v;
        }
    RN_END
}

// ------------------------------------------------------------------------------------------
// activations (rnn.c:41-72)
// ------------------------------------------------------------------------------------------
// The reference goes through double in two places; both are replaced by float operations that give the same bits:
//  * floor((double)(.5f + 25*x)) -- the argument is already a float in [0.5, 200.5] (x in [0, 8) here), its conversion to
//    double is exact and floor of a non-negative value is truncation: (int)(.5f + 25*x).
//  * (float)(.5 + .5*t), t a float in [-1, 1], and (float)(.5*x): halving is exact, so the only rounding is that of
//    .5 + .5t.  For |.5t| >= 2^-30 the exact sum fits in 53 bits (the double is exact, one rounding either way); below that
//    both roundings give exactly .5f.  Hence .5f + .5f*t, and .5f*x, in float.
// tests/test_kernel_emulation.py holds the emulation (this code) against the oracle, which keeps the double form.
RN_HD float tansig_approx(const float *table, float x) {
    float y, dy, sign = 1;
    if (!(x < 8)) return 1;
    if (!(x > -8)) return -1;
    if (x != x) return 0;
    if (x < 0) { x = -x; sign = -1; }
    const int i = (int)(.5f + 25 * x);
    x -= .04f * i;
    y = table[i];
    dy = 1 - y * y;
    y = y + x * dy * (1 - y * x);
    return sign * y;
}
RN_HD float sigmoid_approx(const float *table, float x) {
    return .5f + .5f * tansig_approx(table, .5f * x);
}


// ------------------------------------------------------------------------------------------
// compute_rnn (rnn.c:160-178), executed by a CTA of kRnnThreads threads for R states at once.
// Weights: per layer one matrix W[j][col] (bf16 bit patterns: every int8 weight is exact in bf16), j running
// over the layer's inputs and then, for a GRU, its recurrent inputs -- the order in which compute_gru accumulates
// (rnn.c:122-127).  Columns: z | r | h gates.  Rows are stored in PAIRS: word W2[j/2][col] = W[j][col] | W[j+1][col]<<16
// (every chain length is even), so one 32-bit load feeds two steps of the chain.
// A thread owns one column and runs the chains of all R states interleaved: the weight fetch and unpack are shared
// by the R states, and R independent FADD chains fill each other's latency slots.  Every chain is still evaluated
// in the reference's order on one thread.
// ------------------------------------------------------------------------------------------
// A CTA may hold several independent GROUPS of kRnnThreads threads (each walking its own states, sharing the weights
// in shared memory): phases end at a barrier of the group only (named barrier 1 + group), so the groups drift apart
// and one group's thin phases (a layer narrower than the group) are filled by the other groups' wide ones.
#if defined(__CUDA_ARCH__)
#define RN_CTA_PHASE { const int tid = (int)(threadIdx.x % ::fvad::rn::kRnnThreads);
#define RN_CTA_END } ::fvad::rn::rnn_group_sync();
#else
#define RN_CTA_PHASE for (int tid = 0; tid < ::fvad::rn::kRnnThreads; tid++) {
#define RN_CTA_END }
#endif

constexpr int kRnnThreads = 288;
constexpr int kRnnMaxR = 4;
#if defined(__CUDACC__)
__device__ __forceinline__ void rnn_group_sync() {
#if defined(__CUDA_ARCH__)
    // literal barrier numbers: with a register operand ptxas reserves all 16 barriers for the CTA
    const unsigned g = threadIdx.x / kRnnThreads;
    if (g == 0) asm volatile("bar.sync 1, %0;" ::"n"(kRnnThreads) : "memory");
    else if (g == 1) asm volatile("bar.sync 2, %0;" ::"n"(kRnnThreads) : "memory");
    else asm volatile("bar.sync 3, %0;" ::"n"(kRnnThreads) : "memory");
#endif
}
#endif
struct RnnLayer { int J, M, cols, w_off, b_off; };        // J = chain length, M = inputs before the recurrent part
constexpr int kL_in = 0, kL_vad = 1, kL_vout = 2, kL_noise = 3, kL_den = 4, kL_out = 5;
__host__ __device__ constexpr RnnLayer rnn_layer(int l) {
    return l == 0 ? RnnLayer{42, 42, 24, 0, 0}
         : l == 1 ? RnnLayer{48, 24, 72, 1008, 24}
         : l == 2 ? RnnLayer{24, 24, 1, 4464, 96}
         : l == 3 ? RnnLayer{138, 90, 144, 4488, 97}
         : l == 4 ? RnnLayer{210, 114, 288, 24360, 241}
         :          RnnLayer{96, 96, 22, 84840, 529};
}
constexpr int kRnnWeights = 86952, kRnnBiases = 551;
// vectors of ONE state (floats); state r of a CTA lives at scr + r * kRv_end
constexpr int kRv_feat = 0, kRv_v1 = 48, kRv_v3 = 96, kRv_v4 = 240, kRv_zr = 456, kRv_hp = 648, kRv_hn = 744, kRv_end = 840;

RN_HD unsigned f2u_early(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; unsigned u; } c; c.f = f; return c.u;
#endif
}
RN_HD float u2f_early(unsigned u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { unsigned u; float f; } c; c.u = u; return c.f;
#endif
}
RN_HD float bits_to_f32(unsigned u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { unsigned u; float f; } c; c.u = u; return c.f;
#endif
}
struct RnnF2 { float x, y; };                              // two consecutive vector elements (8-byte aligned: j is even)

// sum[r] += W[j][col]*v_r[j] for j in [j0,j1), in order, for the R states.  j0 is even and j1-j0 a multiple of 6 (true
// for every layer): the chain is walked in blocks of three row pairs, and the next block's weight words and vector
// elements are fetched before the current block is accumulated, so no shared-memory latency sits on the FADD chain.
// EXPERIMENT ONLY (knob rnn_tolerance_mode, never the default): what a tensor-core formulation of the
// gate contractions would do to the numbers, without building one.  An MMA does not keep the reference's summation order
// and takes its activations in reduced precision; the int8 weights are exact in every MMA input type.
//   0  exact: in-order f32 accumulation of f32 products (the product path; modes 1-3 live in an instantiation of their own,
//      k_rn_rnn<1, 1, true>: a run-time switch inside the product kernel's chains cost it 24 % at 592 streams)
//   1  f32 products, accumulated in FOUR interleaved partial sums (any GEMM's freedom to reassociate)
//   2  as 1, activations rounded to TF32 (10-bit mantissa) before the product
//   3  as 1, activations rounded to BF16 (7-bit mantissa)
#if defined(__CUDACC__)
__constant__ int c_rnn_tolerance_mode;
#endif
RN_HD int rnn_tolerance_mode() {
#if defined(__CUDA_ARCH__)
    return c_rnn_tolerance_mode;
#else
    return 0;
#endif
}
RN_HD float rnn_round_act(float x, int mode) {
    if (mode < 2) return x;
    const unsigned u = f2u_early(x);
    const unsigned drop = mode == 2 ? 13u : 16u;             // tf32 keeps 10 mantissa bits, bf16 7: round to nearest even
    const unsigned r = u + ((1u << (drop - 1)) - 1u) + ((u >> drop) & 1u);
    return u2f_early(r & ~((1u << drop) - 1u));
}

constexpr int kRnnP = 3;
template <int R> struct RnnBlk { RnnF2 x[R][kRnnP]; };
struct RnnWBlk { unsigned u[kRnnP]; };
RN_HD void rnn_w_load(RnnWBlk &b, const unsigned *W2, int cols, int col, int j) {
#pragma unroll
    for (int p = 0; p < kRnnP; p++) b.u[p] = W2[((j >> 1) + p) * cols + col];
}
template <int R>
RN_HD void rnn_blk_load(RnnBlk<R> &b, const float *v, int j) {
#pragma unroll
    for (int p = 0; p < kRnnP; p++) {
#pragma unroll
        for (int r = 0; r < R; r++) b.x[r][p] = *reinterpret_cast<const RnnF2 *>(v + r * kRv_end + j + 2 * p);
    }
}
// the experiment's chain: four partial sums, optionally reduced-precision activations (see rnn_tolerance_mode)
template <int R>
RN_HD void rnn_chain_tol(const unsigned *W2, int cols, int col, const float *v, const float *rg, int j0, int j1, float (&sum)[R], int mode) {
    float part[R][4];
    for (int r = 0; r < R; r++) for (int q = 0; q < 4; q++) part[r][q] = 0.f;
    for (int j = j0; j < j1; j += 2) {
        const unsigned u = W2[(j >> 1) * cols + col];
        const float w0 = bits_to_f32(u << 16), w1 = bits_to_f32(u & 0xffff0000u);
        for (int r = 0; r < R; r++) {
            float x0 = v[r * kRv_end + j], x1 = v[r * kRv_end + j + 1];
            if (rg) { x0 *= rg[r * kRv_end + j]; x1 *= rg[r * kRv_end + j + 1]; }     // the reset gate applied to the state first
            part[r][(j >> 1) & 1 ? 2 : 0] += w0 * rnn_round_act(x0, mode);
            part[r][(j >> 1) & 1 ? 3 : 1] += w1 * rnn_round_act(x1, mode);
        }
    }
    for (int r = 0; r < R; r++) sum[r] += (part[r][0] + part[r][1]) + (part[r][2] + part[r][3]);
}

template <int R, bool TOL = false>
RN_HD void rnn_chain(const unsigned *W2, int cols, int col, const float *v, int j0, int j1, float (&sum)[R]) {
    if (TOL) { rnn_chain_tol<R>(W2, cols, col, v, nullptr, j0, j1, sum, rnn_tolerance_mode()); return; }   // the experiment's instantiation only
    RnnBlk<R> cur, nxt;
    RnnWBlk wc, wn;
    rnn_w_load(wc, W2, cols, col, j0);
    rnn_blk_load<R>(cur, v, j0);
    for (int j = j0; j < j1; j += 2 * kRnnP) {
        const int jn = j + 2 * kRnnP < j1 ? j + 2 * kRnnP : j;      // last block: a harmless re-fetch instead of a branch
        rnn_w_load(wn, W2, cols, col, jn);
        rnn_blk_load<R>(nxt, v, jn);
#pragma unroll
        for (int p = 0; p < kRnnP; p++) {
            const float w0 = bits_to_f32(wc.u[p] << 16), w1 = bits_to_f32(wc.u[p] & 0xffff0000u);
#pragma unroll
            for (int r = 0; r < R; r++) {
                sum[r] += w0 * cur.x[r][p].x;
                sum[r] += w1 * cur.x[r][p].y;
            }
        }
        cur = nxt; wc = wn;
    }
}
// recurrent part of the candidate gate: sum += (W[j][col]*state[j-M])*r[j-M]  (rnn.c:147); same blocking
template <int R, bool TOL = false>
RN_HD void rnn_chain_r(const unsigned *W2, int cols, int col, const float *v, const float *rg, int M, int J, float (&sum)[R]) {
    if (TOL) { rnn_chain_tol<R>(W2, cols, col, v, rg - M, M, J, sum, rnn_tolerance_mode()); return; }
    RnnBlk<R> cur, nxt, gcur, gnxt;                         // g*: the reset-gate values
    RnnWBlk wc, wn;
    rnn_w_load(wc, W2, cols, col, M);
    rnn_blk_load<R>(cur, v, M);
    rnn_blk_load<R>(gcur, rg - M, M);
    for (int j = M; j < J; j += 2 * kRnnP) {
        const int jn = j + 2 * kRnnP < J ? j + 2 * kRnnP : j;
        rnn_w_load(wn, W2, cols, col, jn);
        rnn_blk_load<R>(nxt, v, jn);
        rnn_blk_load<R>(gnxt, rg - M, jn);
#pragma unroll
        for (int p = 0; p < kRnnP; p++) {
            const float w0 = bits_to_f32(wc.u[p] << 16), w1 = bits_to_f32(wc.u[p] & 0xffff0000u);
#pragma unroll
            for (int r = 0; r < R; r++) {
                sum[r] += w0 * cur.x[r][p].x * gcur.x[r][p].x;
                sum[r] += w1 * cur.x[r][p].y * gcur.x[r][p].y;
            }
        }
        cur = nxt; gcur = gnxt; wc = wn;
    }
}

constexpr float kWScale = 1.f / 256;

// One GRU layer: v = [inputs(M) | state(N)], zr/hp/hn scratch (all of state 0; state r at + r*kRv_end).  Two CTA phases.
template <int R, bool TOL = false>
RN_HD void rnn_gru_cta(const unsigned *W2, const float *bias, const float *tansig, int layer, float *v, float *zr, float *hp,
                       float *hn) {
    const RnnLayer L = rnn_layer(layer);
    const int N = L.cols / 3;
    const unsigned *Wl = W2 + L.w_off / 2;
    const float *bl = bias + L.b_off;
    RN_CTA_PHASE
        if (tid < 3 * N) {
            float s[R];
#pragma unroll
            for (int r = 0; r < R; r++) s[r] = bl[tid];
            if (tid < 2 * N) {                               // update and reset gates (rnn.c:119-139)
                rnn_chain<R, TOL>(Wl, L.cols, tid, v, 0, L.J, s);
#pragma unroll
                for (int r = 0; r < R; r++) zr[r * kRv_end + tid] = sigmoid_approx(tansig, kWScale * s[r]);
            } else {                                         // candidate: input part (rnn.c:143-145)
                rnn_chain<R, TOL>(Wl, L.cols, tid, v, 0, L.M, s);
#pragma unroll
                for (int r = 0; r < R; r++) hp[r * kRv_end + tid - 2 * N] = s[r];
            }
        }
    RN_CTA_END
    RN_CTA_PHASE
        if (tid < N) {                                       // candidate: recurrent part, ReLU, blend (rnn.c:146-152)
            float s[R];
#pragma unroll
            for (int r = 0; r < R; r++) s[r] = hp[r * kRv_end + tid];
            rnn_chain_r<R, TOL>(Wl, L.cols, 2 * N + tid, v, zr + N, L.M, L.J, s);
#pragma unroll
            for (int r = 0; r < R; r++) {
                float x = kWScale * s[r];
                x = x < 0 ? 0 : x;
                const float z = zr[r * kRv_end + tid];
                hn[r * kRv_end + tid] = z * v[r * kRv_end + L.M + tid] + (1 - z) * x;
            }
        }
    RN_CTA_END
}

template <int R> struct RnnOut { float *gains[R]; float *vad[R]; };

// Whole network for one frame of each of the R states.  scr = vectors (kRv_* offsets) of state 0, state r at
// + r*kRv_end; the frame's features are already at kRv_feat..+42.  Bit r of `act` says state r has a frame this
// round: an inactive state is computed on stale inputs (harmless) but neither its GRU states nor any output is written.
template <int R, bool TOL = false>
RN_HD void rnn_frame_cta(const unsigned *W2, const float *bias, const float *tansig, float *scr, unsigned act, const RnnOut<R> &out) {
    float *feat = scr + kRv_feat, *v1 = scr + kRv_v1, *v3 = scr + kRv_v3, *v4 = scr + kRv_v4;
    float *zr = scr + kRv_zr, *hp = scr + kRv_hp, *hn = scr + kRv_hn;
    RN_CTA_PHASE
        if (tid < 24) {                                      // input_dense, tanh (rnn.c:79-107)
            const RnnLayer L = rnn_layer(kL_in);
            float s[R];
#pragma unroll
            for (int r = 0; r < R; r++) s[r] = bias[L.b_off + tid];
            rnn_chain<R, TOL>(W2 + L.w_off / 2, L.cols, tid, feat, 0, L.J, s);
#pragma unroll
            for (int r = 0; r < R; r++) v1[r * kRv_end + tid] = tansig_approx(tansig, kWScale * s[r]);
        }
    RN_CTA_END
    rnn_gru_cta<R, TOL>(W2, bias, tansig, kL_vad, v1, zr, hp, hn);
    RN_CTA_PHASE
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int o = r * kRv_end;
            if (tid < 24) {                                                                  // vad_gru_state
                if ((act >> r) & 1) v1[o + 24 + tid] = hn[o + tid];
                v3[o + 24 + tid] = hn[o + tid]; v4[o + tid] = hn[o + tid];
            } else if (tid < 48) v3[o + tid - 24] = v1[o + tid - 24];                        // dense_out
            else if (tid < 90) { v3[o + tid] = feat[o + tid - 48]; v4[o + 24 + tid] = feat[o + tid - 48]; }
        }
        if (tid == kRnnThreads - 1) {                        // vad_output, sigmoid (rnn.c:166), on another warp
            const RnnLayer L = rnn_layer(kL_vout);
            float s[R];
#pragma unroll
            for (int r = 0; r < R; r++) s[r] = bias[L.b_off];
            // its input is the NEW vad state, read from hn (v1[24..] is being written by warp 0)
            rnn_chain<R, TOL>(W2 + L.w_off / 2, L.cols, 0, hn, 0, L.J, s);
#pragma unroll
            for (int r = 0; r < R; r++)
                if ((act >> r) & 1) *out.vad[r] = sigmoid_approx(tansig, kWScale * s[r]);
        }
    RN_CTA_END
    rnn_gru_cta<R, TOL>(W2, bias, tansig, kL_noise, v3, zr, hp, hn);
    RN_CTA_PHASE
        if (tid < 48) {                                                                      // noise_gru_state
#pragma unroll
            for (int r = 0; r < R; r++) {
                const int o = r * kRv_end;
                if ((act >> r) & 1) v3[o + 90 + tid] = hn[o + tid];
                v4[o + 24 + tid] = hn[o + tid];
            }
        }
    RN_CTA_END
    rnn_gru_cta<R, TOL>(W2, bias, tansig, kL_den, v4, zr, hp, hn);
    RN_CTA_PHASE
        if (tid < 22) {                                      // denoise_output, sigmoid (rnn.c:177) on the NEW denoise state (hn)
            const RnnLayer L = rnn_layer(kL_out);
            float s[R];
#pragma unroll
            for (int r = 0; r < R; r++) s[r] = bias[L.b_off + tid];
            rnn_chain<R, TOL>(W2 + L.w_off / 2, L.cols, tid, hn, 0, L.J, s);
#pragma unroll
            for (int r = 0; r < R; r++)
                if ((act >> r) & 1) out.gains[r][tid] = sigmoid_approx(tansig, kWScale * s[r]);
        } else if (tid >= 32 && tid < 128) {                                                 // denoise_gru_state, warps 1-3
#pragma unroll
            for (int r = 0; r < R; r++)
                if ((act >> r) & 1) v4[r * kRv_end + 114 + tid - 32] = hn[r * kRv_end + tid - 32];
        }
    RN_CTA_END
}

// biquad (denoise.c:410-420) with the Denoiser's x32767 input scaling (Denoiser.zig:72-82)
// xi = the sample already scaled by 32767 (Denoiser.zig:45-66 normalizedPcmToRnnoise)
RN_HD float biquad_step_x(float xi, float &m0, float &m1) {
    const float a0 = -1.99599f, a1 = 0.99600f, b0 = -2.f, b1 = 1.f;     // denoise.c:470-471
    const float yi = xi + m0;
    // a0*yi and a1*yi are products of two floats: exact in double, so DMUL+DADD == one DFMA bit for bit
    // (b0*xi = -2*xi and b1*xi = xi are exact as well).  Shortens the dependent chain by one DADD.
    const double xd = (double)xi, yd = (double)yi;
    (void)b0; (void)b1;
    m0 = (float)((double)m1 + fma(-(double)a0, yd, -2.0 * xd));
    m1 = (float)fma(-(double)a1, yd, xd);
    return yi;
}
RN_HD float biquad_step(float in_norm, float &m0, float &m1) { return biquad_step_x(in_norm * 32767.0f, m0, m1); }

// ------------------------------------------------------------------------------------------
// The same biquad step WITHOUT double precision: error-free transformations in f32 (TwoSum, FMA
// products) evaluate E0 = m1 - 2x + A0*y and E1 = x - A1*y exactly as short expansions and round them
// once.  The reference rounds twice (to double, then to float) -- the two agree unless the exact
// value lies within ~2^-29 ulp of a rounding boundary.  Each step therefore also computes, OFF the
// dependent chain, the exact distance of E from the rounding boundary and raises `hazard` when it is
// within 2^-14 ulp of one, or when a tiny result, an underflowing product or a cancellation beyond 2^12
// (with an inexact double intermediate) makes the bound unsafe.  A hazard never changes the result: the caller discards the
// block and replays it with biquad_step (the reference's own double arithmetic).  Dependent chain per
// sample: FADD, FMUL, FADD, TwoSum tail (4), FADD, FADD = 9 f32 ops (~38 cycles) instead of
// FADD, F2F, DFMA, DADD, F2F (~84 cycles measured on B200).
// ------------------------------------------------------------------------------------------
RN_HD unsigned f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; unsigned u; } c; c.f = f; return c.u;
#endif
}
RN_HD float u2f(unsigned u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; unsigned u; } c; c.u = u; return c.f;
#endif
}
RN_HD float fma_rn(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
RN_HD void two_sum(float a, float b, float &s, float &e) {
    s = a + b;
    const float bb = s - a;
    e = (a - (s - bb)) + (b - bb);
}
// r = RN(w + uh) was computed with exact residual v (E = r + v up to ~2^-46 |v|): is r = RN(E) beyond doubt?
// `big` bounds the magnitudes whose 2^-53 relative perturbation the reference's double roundings add.
RN_HD bool round_hazard(float r, float v, float big) {
    const unsigned ur = f2u(r), e = (ur >> 23) & 0xffu;
    const float av = v < 0 ? -v : v, ar = r < 0 ? -r : r;
    if (e <= 26u || e >= 250u) return !(ar == 0.0f && av == 0.0f && big == 0.0f);   // tiny / zero / huge results: not analysed
    float half = u2f((e - 24u) << 23);                         // ulp(r) / 2
    if ((ur & 0x7fffffu) == 0u) half *= 0.5f;                  // below a power of two the grid is twice as fine
    return (av >= half * 0.99993896484375f) || (big >= ar * 4096.0f);
}
RN_HD float biquad_step_eft_x(float x, float &m0, float &m1, bool &hazard) {     // x already scaled by 32767
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float c = -2.0f * x;                                 // exact
    float dh, dl;
    two_sum(m1, c, dh, dl);                                    // m1 - 2x = dh + dl          (off the chain)
    const float y = x + m0;                                    // the reference's own float add
    // ---- m0' = RN(m1 - 2x + A0*y) -------------------------------------------------------------
    const float p = A0 * y, pl = fma_rn(A0, y, -p);            // A0*y = p + pl
    float w, f;
    two_sum(p, dh, w, f);                                      // p + dh = w + f
    float th, tl, uh, ul, r, g;
    two_sum(pl, dl, th, tl);
    two_sum(th, f, uh, ul);                                    // pl + dl + f = uh + ul + tl
    two_sum(w, uh, r, g);                                      // E0 = r + g + ul + tl
    // A0*y - 2x and x - A1*y are EXACT in double when x == 0 or the exponents of x and y are within 2
    // (48-bit product + 24-bit term span <= 53 bits): then the reference's inner rounding perturbs nothing.
    const int ex = (int)((f2u(x) >> 23) & 0xffu), ey = (int)((f2u(y) >> 23) & 0xffu);
    const bool inner_exact = (x == 0.0f) || (ex - ey <= 2 && ey - ex <= 2);
    const float ap = p < 0 ? -p : p, ac = c < 0 ? -c : c;
    bool hz = round_hazard(r, g + (ul + tl), inner_exact ? 0.0f : ap + ac);
    hz |= (ap != 0.0f && ap < 1.0e-30f);                       // fma error term would underflow
    // ---- m1' = RN(x - A1*y) ---------------------------------------------------------------------
    const float q = A1 * y, ql = fma_rn(A1, y, -q);            // A1*y = q + ql
    float s, e, u, u2, n1, g1;
    two_sum(x, -q, s, e);                                      // x - q = s + e
    two_sum(e, -ql, u, u2);
    two_sum(s, u, n1, g1);                                     // E1 = n1 + g1 + u2
    const float aq = q < 0 ? -q : q, ax = x < 0 ? -x : x;
    hz |= round_hazard(n1, g1 + u2, inner_exact ? 0.0f : aq + ax);
    hazard |= hz;
    m0 = r;
    m1 = n1;
    return y;
}
RN_HD float biquad_step_eft(float in_norm, float &m0, float &m1, bool &hazard) {
    return biquad_step_eft_x(in_norm * 32767.0f, m0, m1, hazard);
}

// The chain lane's step, second form (k_rn_biquad_v2): the sample arrives scaled, nothing but the state is produced
// (the verifying threads form y = x + m0 themselves, it is the reference's own float add), and -2x is never
// materialised: RN(m1 - 2x) and RN(-2x - bb) are single fused operations on the exact product, bit-identical to the
// add / subtract of the exact float -2x.  20 instructions per sample instead of 23, the same 7 on the dependent chain.
RN_HD void biquad_chain_step(float x, float &m0, float &m1) {
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float dh = fma_rn(-2.0f, x, m1);                     // TwoSum(m1, -2x): dh + dl = m1 - 2x
    const float bb = dh - m1;
    const float dl = (m1 - (dh - bb)) + fma_rn(-2.0f, x, -bb);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    const float w = p + dh;
    const bool pbig = fabsf(y) >= fabsf(dh) * (1.0f / 1.99599f);       // ordering decided beside p and w (see biquad_step_fast)
    const float hi = pbig ? p : dh, lo = pbig ? dh : p;
    const float f = (hi - w) + lo;
    const float th = pl + dl;
    const float uh = th + f;
    m1 = fma_rn(-A1, y, x);
    m0 = w + uh;
}

// The chain-only version of biquad_step_eft (same operations for y, m0', m1'; no hazard analysis): what the
// one sequential lane executes.  ~30 instructions per sample, 9 of them on the dependent chain.
RN_HD float biquad_step_fast(float in_norm, float &m0, float &m1) {
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float x = in_norm * 32767.0f;
    const float c = -2.0f * x;
    float dh, dl;
    two_sum(m1, c, dh, dl);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    // error term of w = p + dh by Fast2Sum on the magnitude-ordered pair (2 dependent operations after w
    // instead of TwoSum's 4; the ordering is decided while w is being computed)
    // The ordering is decided from y, not from p = A0*y: |y| >= |dh|/A0 is known one multiply earlier, so the compare and
    // the two selects run beside p and w instead of between w and (hi - w) -- one operation less on the dependent chain.
    // Near |p| == |dh| the two tests can disagree; Fast2Sum is then still exact unless the operands straddle a binade,
    // and ANY difference from the verified step (biquad_step_eft uses the branch-free TwoSum) is caught by the verifying
    // warp's comparison and replayed -- the fast step only has to be right almost always.
    const float w = p + dh;
    const bool pbig = fabsf(y) >= fabsf(dh) * (1.0f / 1.99599f);
    const float hi = pbig ? p : dh, lo = pbig ? dh : p;
    const float f = (hi - w) + lo;
    const float th = pl + dl;
    const float uh = th + f;
    const float r = w + uh;
    m1 = fma_rn(-A1, y, x);                                    // RN(x - A1*y): one rounding of the exact value
    m0 = r;
    return y;
}

// One block of 32 samples for one state, executed by a warp: lane 0 runs the chain with the fast step and
// records the state before every sample; then all 32 lanes re-run their sample with the full hazard
// analysis in parallel and compare their result with the state lane 0 carried on; if any lane objects,
// lane 0 replays the block with the reference's double arithmetic.  (m0, m1) is the filter state, meaningful
// on lane 0 (a register on the device); sm0/sm1 = 33 floats each; flag = 1 int (0 on entry and on exit).
// While the state is tiny (digital silence decaying into denormals) every block would be hazardous, so the
// block goes straight to the double path.
RN_HD void biquad_block_warp(const float *xin, float *yout, float &m0, float &m1, float *sm0, float *sm1, int *flag) {
    float e0 = 0, e1 = 0;                                   // lane 0: state at the end of the fast pass
    RN_PHASE
        if (lane == 0) {
            const float a0 = m0 < 0 ? -m0 : m0, a1 = m1 < 0 ? -m1 : m1;
            if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
                for (int i = 0; i < 32; i++) yout[i] = biquad_step(xin[i], m0, m1);
                *flag = 2;                                  // already final
            } else {
                float xr[32];                               // all loads first: nothing on the chain waits for memory
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = xin[i];
                e0 = m0; e1 = m1;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    sm0[i] = e0; sm1[i] = e1;
                    yout[i] = biquad_step_fast(xr[i], e0, e1);
                }
                sm0[32] = e0; sm1[32] = e1;
            }
        }
    RN_END
    RN_PHASE
        if (*flag != 2) {
            float a = sm0[lane], b = sm1[lane];
            bool hz = false;
            const float yv = biquad_step_eft(xin[lane], a, b, hz);
            hz |= (f2u(a) != f2u(sm0[lane + 1])) | (f2u(b) != f2u(sm1[lane + 1])) | (f2u(yv) != f2u(yout[lane]));
            if (hz) *flag = 1;                              // any lane may set it; same value
        }
    RN_END
    RN_PHASE
        if (lane == 0) {
            const int fl = *flag;
            if (fl == 1) {
                for (int i = 0; i < 32; i++) yout[i] = biquad_step(xin[i], m0, m1);
            } else if (fl == 0) {
                m0 = e0; m1 = e1;
            }
            if (fl) *flag = 0;
        }
    RN_END
}

}  // namespace rn
}  // namespace fvad
