// rnnoise.cu -- batched RNNoise on the GPU: kernels + tile orchestration (see rnnoise.cuh).
//
// What is sequential in rnnoise_process_frame (denoise.c:458-494) and what is not:
//   sample-sequential : biquad (denoise.c:410-420)                          -> k_rn_biquad, one lane per state
//   frame-sequential  : remove_doubling's prev_period/prev_gain decision      -> k_rn_pitch_seq (tiny)
//                       cepstral ring position (only non-silent frames count) -> k_rn_compact (prefix count)
//                       the three GRUs (rnn.c:109-156)                        -> k_rn_rnn, one CTA per state
//                       lastg floor (denoise.c:478-482)                       -> k_rn_gain_scan
//   frame-parallel    : everything else (3 FFTs, band energies, pitch correlations, DCTs, synthesis)
//                                                                             -> warp-per-frame kernels
// All arithmetic lives in rn_core.cuh and runs in the reference's operation order.
#include "rnnoise.cuh"
#include "rn_core.cuh"
#include "rn_tables.cuh"
#include <algorithm>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

namespace fvad {
using namespace rn;

struct RnPersist {                 // per state, lives across tiles (DenoiseState, denoise.c:86-98)
    float biq[2];
    int last_period;
    float last_gain;
    long long nz_total;            // non-silent frames so far: cepstral ring position = nz_total & 7
    float lastg[kNB];
    float rnn_vec[kRv_end];        // RNN vectors incl. the three GRU states
};

struct RnDevTables {
    Tables t{};
    unsigned *weights2 = nullptr;            // row-paired bf16 matrices (rn_core.cuh)
    float *bias = nullptr;
};

// Tile arrays that cross a stage boundary exist once per SLOT so that stage X of tile t+1 can run while
// stage X+1 of tile t still reads tile t's arrays (batch.cu runs the stages on CUDA streams of their own).
struct RnSlot {
    float *hp = nullptr;           // [N][kPHist + Qcap*480]   A -> B
    cpx *X = nullptr, *P = nullptr;          // [N][Qcap][481]  B -> D
    float *Ex = nullptr, *Ep = nullptr, *Exp = nullptr, *ceps = nullptr, *gains = nullptr, *gsm = nullptr;   // [N][Qcap][22]
    float *featB = nullptr;        // [N][Qcap][8]
    float *feat = nullptr;         // [N][Qcap][42]            B -> C
    float *vad = nullptr;          // [N][Qcap]                C -> D
    int *sil = nullptr, *pidx = nullptr, *nzlist = nullptr;   // [N][Qcap]
    int *nz_tile = nullptr;        // [N] non-silent frames of the tile
    long long *rank = nullptr;     // [N][Qcap]
    PitchCand *cand = nullptr;     // [N][Qcap]
};

struct RnnoiseBatch {
    int device = 0;
    size_t N = 0, Qcap = 0;
    RnDevTables tab;
    RnPersist *persist = nullptr;
    std::vector<RnSlot> slots;
    // single-stage arrays (carried in place)
    float *hp_hist = nullptr;      // [N][kPHist]   stage A: pitch/analysis history between tiles
    // stage A, fourth biquad form: what the chain kernel publishes for the verifying kernel (one tile; stage A of the next
    // tile follows on the same stream)
    float *bq_m0 = nullptr;        // [N][bq_m0_stride()]  m0 before every sample of the tile, then the end state's
    float *bq_m1 = nullptr;        // [N][bq_m1_stride()]  m1 before every 8th sample, then the end state's
    int *bq_broken = nullptr;      // [N]  first block whose replay disagreed with the chain (kBq4None: none)
    size_t bq_m0_stride() const { return Qcap * kFrame + 4; }
    size_t bq_m1_stride() const { return Qcap * (kFrame / 8) + 4; }
    float *cepsc = nullptr;        // [N][8 + Qcap][22]  stage B
    float *head = nullptr;         // [N][Qcap][480]     stage D
    float *tail = nullptr;         // [N][1 + Qcap][480] stage D
    std::vector<void *> allocs;
    size_t hp_stride() const { return kPHist + Qcap * kFrame; }
};

// ------------------------------------------------------------------------------------------
// tables per device
// ------------------------------------------------------------------------------------------
static std::mutex g_rn_mu;
static RnDevTables g_rn_tab[64];
static bool g_rn_ready[64];

template <class T> static int up(T **dst, const std::vector<T> &src) {
    FVAD_CUDA_TRY(cudaMalloc((void **)dst, sizeof(T) * src.size()));
    FVAD_CUDA_TRY(cudaMemcpy(*dst, src.data(), sizeof(T) * src.size(), cudaMemcpyHostToDevice));
    return FVAD_OK;
}

static int rn_tables_get(int device, RnDevTables *out) {
    std::lock_guard<std::mutex> lk(g_rn_mu);
    if (!g_rn_ready[device]) {
        HostTables h;
        FVAD_TRY(host_tables_build(&h, nullptr));
        RnDevTables d;
        float *hw, *dct, *frac, *omf, *ts; cpx *tw; short *rev;
        FVAD_TRY(up(&hw, h.half_window)); FVAD_TRY(up(&dct, h.dct)); FVAD_TRY(up(&frac, h.frac)); FVAD_TRY(up(&omf, h.omf));
        FVAD_TRY(up(&ts, h.tansig)); FVAD_TRY(up(&tw, h.tw)); FVAD_TRY(up(&rev, h.rev));
        FVAD_TRY(up(&d.weights2, h.weights2)); FVAD_TRY(up(&d.bias, h.bias));
        d.t.half_window = hw; d.t.dct = dct; d.t.frac = frac; d.t.omf = omf; d.t.tansig = ts; d.t.tw = tw; d.t.rev = rev;
        d.t.dct_scale = h.dct_scale;
        g_rn_tab[device] = d;
        g_rn_ready[device] = true;
    }
    *out = g_rn_tab[device];
    return FVAD_OK;
}

// ------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------
// One chain lane + verifying warps per state.  The IIR is sample-sequential, so ONE lane (warp 0, lane 0) walks the
// samples with the fast f32 step (rn_core.cuh biquad_step_fast, 7 dependent f32 ops per sample) and records the state
// before every sample.  Warps 1.. run one block (BQ samples, a sample per thread) behind: they re-evaluate every sample
// of the previous block with the full hazard analysis (biquad_step_eft) and compare with what lane 0 carried on; warp 1
// also stages frames between global and shared memory, so neither verification nor memory latency sits on the
// dependent chain.  A hazardous block is replayed by lane 0 with the reference's double arithmetic from the recorded
// state; the replay reproduces the fast result unless the fast step was really wrong, in which case the block in
// flight is recomputed as well.
// BQ = samples per block, ROLL = the chain lane's loop form.  The per-block costs (two barriers, verdict) shrink with BQ.
// Unrolled form (inputs of two blocks in registers, statically indexed): the body grows with BQ -- measured, 20 streams,
// ms per 60 s tile: BQ = 32: 74.3, 48: 69.6, 60: 77.7, 80: 86.6, and inside the wavefront 48 costs 75.9: beyond the 32 KB
// L1.5 instruction cache a kernel with a small grid is throttled (B300_MICROARCH.md, I-cache), and the other stages'
// kernels compete for the shared instruction cache levels.  Rolled form (16 samples per iteration, ~6 KB body, lives in
// the lowest instruction cache levels): BQ = 48: 71.0, 96: 66.2, 120: 67.8, 160: 66.7 ms alone and 96 stays at 69 ms inside the
// wavefront -- the default.  knob bq_block / knob bq_roll override.
__host__ __device__ constexpr int bq_threads(int BQ) { return 32 + ((BQ + 31) / 32) * 32; }
#define FVAD_BQ_VARIANTS(X) X(32, false) X(48, false) X(60, false) X(48, true) X(96, true) X(160, true)
#ifndef FVAD_BQ_SUB
#define FVAD_BQ_SUB 16      // samples per iteration of the rolled chain loop; ms per tile: 4: 64.1, 8: 63.0, 16: 61.0, 24: 60.6, 32: 61.4, 48: 65.7
#endif
template <int BQ, bool ROLL>
__global__ void __launch_bounds__(bq_threads(BQ)) k_rn_biquad(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                  RnPersist *persist, float *__restrict__ hp, size_t hp_stride, size_t N) {
    constexpr int kBPF = kFrame / BQ, kThreads = bq_threads(BQ);
    static_assert(kFrame % BQ == 0 && BQ % 4 == 0 && (!ROLL || BQ % 8 == 0), "blocks tile the frame");
    const size_t n = blockIdx.x;
    if (n >= N) return;
    __shared__ __align__(16) float sx[2][kFrame];
    __shared__ __align__(16) float sy[2][kFrame];
    __shared__ __align__(16) float2 P[2][BQ + 2];          // (m0, m1) before every sample of the block, + the end state
    __shared__ int FINAL[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float m0 = persist[n].biq[0], m1 = persist[n].biq[1];         // meaningful on thread 0
    float *dst = hp + n * hp_stride + kPHist;
    auto src = [&](int q) { return in + (n * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame; };
    if (threadIdx.x < 2) FINAL[threadIdx.x] = 0;
    if (Q > 0) for (int i = threadIdx.x; i < kFrame; i += kThreads) sx[0][i] = src(0)[i];
    __syncthreads();
    const int G = Q * kBPF;
    float pf[kFrame / 32]; int pf_buf = -1;                        // warp 1: a frame in flight from global memory
    float xr[ROLL ? 1 : BQ];                                 // thread 0 (unrolled form): the block's input, fetched one block ahead
    if (!ROLL && threadIdx.x == 0 && G > 0) {
#pragma unroll
        for (int i = 0; i < BQ; i++) xr[i] = sx[0][i];
    }
    for (int g = 0; g <= G + 1; g++) {
        bool hz = false;
        if (warp == 0) {
            if (lane == 0 && g < G) {                              // ---- the chain: block g
                const int q = g / kBPF, b0 = (g - q * kBPF) * BQ, par = g & 1;
                float *yout = &sy[q & 1][b0];
                if constexpr (ROLL) {
                    // Rolled form: FVAD_BQ_SUB samples per loop iteration, the next iteration's inputs fetched (16-byte loads)
                    // before this iteration's stores queue up.  The loop body (~6 KB) stays in the lowest instruction cache
                    // levels whatever BQ is, so the block can be long (fewer barriers and verdicts per sample) without the
                    // unrolled form's code size.
                    const float *xin = &sx[q & 1][b0];
                    const float a0 = fabsf(m0), a1 = fabsf(m1);
                    if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
                        for (int i = 0; i < BQ; i++) yout[i] = biquad_step(xin[i], m0, m1);   // denormal tail of a silence
                        FINAL[par] = 1;
                    } else {
                        constexpr int kSub = BQ % FVAD_BQ_SUB == 0 ? FVAD_BQ_SUB : 8;   // samples per iteration
                        float4 xc[kSub / 4], xn[kSub / 4];
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xc[k] = *reinterpret_cast<const float4 *>(xin + 4 * k);
#pragma unroll 1
                        for (int i = 0; i < BQ; i += kSub) {
                            const int inx = i + kSub < BQ ? i + kSub : i;     // last iteration: a harmless re-fetch instead of a branch
#pragma unroll
                            for (int k = 0; k < kSub / 4; k++) xn[k] = *reinterpret_cast<const float4 *>(xin + inx + 4 * k);
#pragma unroll
                            for (int k = 0; k < kSub / 4; k++) {
                                float4 sa, sb, yy;
                                sa.x = m0; sa.y = m1; yy.x = biquad_step_fast(xc[k].x, m0, m1);
                                sa.z = m0; sa.w = m1; yy.y = biquad_step_fast(xc[k].y, m0, m1);
                                sb.x = m0; sb.y = m1; yy.z = biquad_step_fast(xc[k].z, m0, m1);
                                sb.z = m0; sb.w = m1; yy.w = biquad_step_fast(xc[k].w, m0, m1);
                                *reinterpret_cast<float4 *>(&P[par][i + 4 * k]) = sa;
                                *reinterpret_cast<float4 *>(&P[par][i + 4 * k + 2]) = sb;
                                *reinterpret_cast<float4 *>(&yout[i + 4 * k]) = yy;
                            }
#pragma unroll
                            for (int k = 0; k < kSub / 4; k++) xc[k] = xn[k];
                        }
                        P[par][BQ] = make_float2(m0, m1);
                        FINAL[par] = 0;
                    }
                } else {
                float xn[BQ];                                // next block's input: loaded before this block's stores queue up
                if (g + 1 < G) {
                    const int q1 = (g + 1) / kBPF, b1 = (g + 1 - q1 * kBPF) * BQ;
#pragma unroll
                    for (int i = 0; i < BQ; i++) xn[i] = sx[q1 & 1][b1 + i];
                }
                const float a0 = fabsf(m0), a1 = fabsf(m1);
                if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
#pragma unroll
                    for (int i = 0; i < BQ; i++) yout[i] = biquad_step(xr[i], m0, m1);   // denormal tail of a silence
                    FINAL[par] = 1;
                } else {
                    // one lane issues every shared-memory store of the block: pack 4 samples per store
#pragma unroll
                    for (int i = 0; i < BQ; i += 4) {
                        float4 sa, sb, yy;
                        sa.x = m0; sa.y = m1; yy.x = biquad_step_fast(xr[i], m0, m1);
                        sa.z = m0; sa.w = m1; yy.y = biquad_step_fast(xr[i + 1], m0, m1);
                        sb.x = m0; sb.y = m1; yy.z = biquad_step_fast(xr[i + 2], m0, m1);
                        sb.z = m0; sb.w = m1; yy.w = biquad_step_fast(xr[i + 3], m0, m1);
                        *reinterpret_cast<float4 *>(&P[par][i]) = sa;
                        *reinterpret_cast<float4 *>(&P[par][i + 2]) = sb;
                        *reinterpret_cast<float4 *>(&yout[i]) = yy;
                    }
                    P[par][BQ] = make_float2(m0, m1);
                    FINAL[par] = 0;
                }
#pragma unroll
                for (int i = 0; i < BQ; i++) xr[i] = xn[i];
                }
            }
        } else {
            const int v = (int)threadIdx.x - 32;                   // verification: one sample of the block per thread of warps >= 1
            if (warp == 1 && pf_buf >= 0) {                        // ---- the frame fetched one iteration ago lands in shared memory
#pragma unroll
                for (int k = 0; k < kFrame / 32; k++) sx[pf_buf][lane + 32 * k] = pf[k];
                pf_buf = -1;
            }
            if (g >= 1 && g - 1 < G && v < BQ) {             // ---- verification of block g-1
                const int gv = g - 1, q = gv / kBPF, b0 = (gv - q * kBPF) * BQ, par = gv & 1;
                if (!FINAL[par]) {
                    const float2 pre = P[par][v], post = P[par][v + 1];
                    float a = pre.x, b = pre.y;
                    const float yv = biquad_step_eft(sx[q & 1][b0 + v], a, b, hz);
                    hz |= (__float_as_uint(a) != __float_as_uint(post.x)) |
                          (__float_as_uint(b) != __float_as_uint(post.y)) |
                          (__float_as_uint(yv) != __float_as_uint(sy[q & 1][b0 + v]));
                }
            }
            if (warp == 1 && g >= 2) {                             // ---- block g-2 is final: frame I/O at frame edges
                const int gf = g - 2, q = gf / kBPF;
                if (gf - q * kBPF == kBPF - 1) {
                    for (int i = lane; i < kFrame; i += 32) dst[(size_t)q * kFrame + i] = sy[q & 1][i];
                    // sx[q & 1] is free once frame q is final: fetch frame q + 2 (frame q + 1 was fetched a frame ago)
                    // (into registers now, into shared memory next iteration: the load latency never meets the barrier)
                    if (q + 2 < Q) {
                        const float *s2 = src(q + 2);
#pragma unroll
                        for (int k = 0; k < kFrame / 32; k++) pf[k] = s2[lane + 32 * k];
                        pf_buf = q & 1;
                    }
                }
            }
            if (warp == 1 && g == 0 && Q > 1) {
                const float *s1 = src(1);
#pragma unroll
                for (int k = 0; k < kFrame / 32; k++) pf[k] = s1[lane + 32 * k];
                pf_buf = 1;
            }
        }
        const int verdict = __syncthreads_or(hz ? 1 : 0);          // the verdict on block g-1 travels with the barrier
        if (threadIdx.x == 0 && verdict) {                         // ---- replay of a hazardous block g-1
            const int gv = g - 1, q = gv / kBPF, b0 = (gv - q * kBPF) * BQ, par = gv & 1;
            float r0 = P[par][0].x, r1 = P[par][0].y;
            for (int i = 0; i < BQ; i++) sy[q & 1][b0 + i] = biquad_step(sx[q & 1][b0 + i], r0, r1);
            if (__float_as_uint(r0) != __float_as_uint(P[par][BQ].x) || __float_as_uint(r1) != __float_as_uint(P[par][BQ].y)) {
                // the fast step was wrong: block g (computed from its end state) is void, redo it from the true state
                m0 = r0; m1 = r1;
                if (g < G) {
                    const int q2 = g / kBPF, c0 = (g - q2 * kBPF) * BQ;
                    for (int i = 0; i < BQ; i++) sy[q2 & 1][c0 + i] = biquad_step(sx[q2 & 1][c0 + i], m0, m1);
                    FINAL[g & 1] = 1;
                }
            }
        }
        __syncthreads();                                           // replay results visible before anyone moves on
    }
    if (threadIdx.x == 0) { persist[n].biq[0] = m0; persist[n].biq[1] = m1; }
}

// Second form of the chain-per-CTA kernel (knob bq_form=2; an experiment of round 2 that did NOT pay: 250 ms per 240 s of 20
// streams against 244 ms for the first form).  The premise was that the chain lane is issue-bound: 42 cycles per sample
// against a 28-cycle dependent chain with ~25 instructions per sample on ONE warp.  tools/ubench/chain_mix.cu says otherwise:
// one warp issues independent FADD / FFMA / FMUL at 1.09 cycles each, the 7-operation chain alone costs 28.4 cycles, the whole
// step in registers 30.8 (the two FSELs that order the Fast2Sum operands sit on the ALU pipe: +1 cycle each way), with its
// shared-memory traffic 33-39, inside the block structure with barriers and verifying warps 36.3 -- the kernel's 41.6 is
// neither instruction count (21 vs 25 per sample changes nothing) nor barriers (0.5 cycles per sample) nor replays (2 %) nor
// the frame I/O (1.4 cycles); a bare chain with everything else switched off (knob bq_dbg=23) still costs 40.3.  What this
// form takes off the lane:
//  * the scaling by 32767 (the staging warp stores x * 32767: every consumer wants the scaled sample) and the product -2x
//    (folded into two FFMAs, rn_core.cuh biquad_chain_step);
//  * every y store and the m1 half of the state record: the lane publishes m0 before each sample (one STS.128 per four
//    samples instead of three) plus m1 at the block's ends.  A verifying thread rebuilds its sample's m1 the way the lane
//    formed it, fma(-A1, x' + m0', x') of the PREVIOUS sample, re-evaluates the step with the full hazard analysis, compares
//    the m0 the lane carried on and its own fused m1 against the analysed results, and stores y = x + m0 itself.
// Protocol, replay and the denormal-tail path are those of k_rn_biquad.
template <int BQ>
__global__ void __launch_bounds__(bq_threads(BQ)) k_rn_biquad_v2(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                                 RnPersist *persist, float *__restrict__ hp, size_t hp_stride, size_t N,
                                                                 int dbg) {
    // dbg (knob bq_dbg, measurement only, results then unverified): 1 = no hazard analysis, 2 = verdicts ignored
    constexpr int kBPF = kFrame / BQ, kThreads = bq_threads(BQ);
    constexpr int kSub = BQ % FVAD_BQ_SUB == 0 ? FVAD_BQ_SUB : 8;   // samples per iteration of the chain loop
    static_assert(kFrame % BQ == 0 && BQ % 8 == 0, "blocks tile the frame");
    const size_t n = blockIdx.x;
    if (n >= N) return;
    __shared__ __align__(16) float sx[2][kFrame];          // input scaled by 32767
    __shared__ __align__(16) float sy[2][kFrame];
    __shared__ __align__(16) float PM[2][BQ + 4];          // m0 before every sample of the block; [BQ] = after the block
    __shared__ float M1[2][2];                             // m1 before / after the block
    __shared__ int FINAL[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float m0 = persist[n].biq[0], m1 = persist[n].biq[1];         // meaningful on thread 0
    float *dst = hp + n * hp_stride + kPHist;
    auto src = [&](int q) { return in + (n * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame; };
    if (threadIdx.x < 2) FINAL[threadIdx.x] = 0;
    if (Q > 0) for (int i = threadIdx.x; i < kFrame; i += kThreads) sx[0][i] = src(0)[i] * 32767.0f;
    __syncthreads();
    const int G = Q * kBPF;
    float pf[kFrame / 32]; int pf_buf = -1;                        // warp 1: a frame in flight from global memory
    for (int g = 0; g <= G + 1; g++) {
        bool hz = false;
        if (warp == 0) {
            if (lane == 0 && g < G) {                              // ---- the chain: block g
                const int q = g / kBPF, b0 = (g - q * kBPF) * BQ, par = g & 1;
                const float *xin = &sx[q & 1][b0];
                const float a0 = fabsf(m0), a1 = fabsf(m1);
                if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
                    float *yout = &sy[q & 1][b0];
                    for (int i = 0; i < BQ; i++) yout[i] = biquad_step_x(xin[i], m0, m1);     // denormal tail of a silence
                    FINAL[par] = 1;
                } else {
                    M1[par][0] = m1;
                    float4 xc[kSub / 4], xn[kSub / 4];
#pragma unroll
                    for (int k = 0; k < kSub / 4; k++) xc[k] = *reinterpret_cast<const float4 *>(xin + 4 * k);
#pragma unroll 1
                    for (int i = 0; i < BQ; i += kSub) {
                        const int inx = i + kSub < BQ ? i + kSub : i;     // last iteration: a harmless re-fetch instead of a branch
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xn[k] = *reinterpret_cast<const float4 *>(xin + inx + 4 * k);
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) {
                            float4 pm;
                            pm.x = m0; biquad_chain_step(xc[k].x, m0, m1);
                            pm.y = m0; biquad_chain_step(xc[k].y, m0, m1);
                            pm.z = m0; biquad_chain_step(xc[k].z, m0, m1);
                            pm.w = m0; biquad_chain_step(xc[k].w, m0, m1);
                            *reinterpret_cast<float4 *>(&PM[par][i + 4 * k]) = pm;
                        }
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xc[k] = xn[k];
                    }
                    PM[par][BQ] = m0;
                    M1[par][1] = m1;
                    FINAL[par] = 0;
                }
            }
        } else {
            const int v = (int)threadIdx.x - 32;                   // verification: one sample of the block per thread of warps >= 1
            if (warp == 1 && pf_buf >= 0) {                        // ---- the frame fetched one iteration ago lands in shared memory
#pragma unroll
                for (int k = 0; k < kFrame / 32; k++) sx[pf_buf][lane + 32 * k] = pf[k] * 32767.0f;
                pf_buf = -1;
            }
            if (g >= 1 && g - 1 < G && v < BQ && !(dbg & 16)) {    // ---- verification of block g-1
                const int gv = g - 1, q = gv / kBPF, b0 = (gv - q * kBPF) * BQ, par = gv & 1;
                if (!FINAL[par]) {
                    const float *xb = &sx[q & 1][b0];
                    float a = PM[par][v], b;
                    if (v == 0) b = M1[par][0];
                    else { const float xp = xb[v - 1]; b = __fmaf_rn(-0.99600f, xp + PM[par][v - 1], xp); }
                    const float xv = xb[v];
                    float yv;
                    if (dbg & 1) yv = xv + a;
                    else {
                        yv = biquad_step_eft_x(xv, a, b, hz);
                        hz |= (__float_as_uint(a) != __float_as_uint(PM[par][v + 1])) |
                              (__float_as_uint(b) != __float_as_uint(__fmaf_rn(-0.99600f, yv, xv)));
                    }
                    if (!(dbg & 8)) sy[q & 1][b0 + v] = yv;        // the reference's own float add: nothing to verify
                }
            }
            if (warp == 1 && g >= 2) {                             // ---- block g-2 is final: frame I/O at frame edges
                const int gf = g - 2, q = gf / kBPF;
                if (gf - q * kBPF == kBPF - 1 && !(dbg & 4)) {
                    for (int i = lane; i < kFrame; i += 32) dst[(size_t)q * kFrame + i] = sy[q & 1][i];
                    if (q + 2 < Q) {
                        const float *s2 = src(q + 2);
#pragma unroll
                        for (int k = 0; k < kFrame / 32; k++) pf[k] = s2[lane + 32 * k];
                        pf_buf = q & 1;
                    }
                }
            }
            if (warp == 1 && g == 0 && Q > 1) {
                const float *s1 = src(1);
#pragma unroll
                for (int k = 0; k < kFrame / 32; k++) pf[k] = s1[lane + 32 * k];
                pf_buf = 1;
            }
        }
        const int verdict = __syncthreads_or(hz ? 1 : 0);          // the verdict on block g-1 travels with the barrier
        if (threadIdx.x == 0 && verdict && !(dbg & 2)) {           // ---- replay of a hazardous block g-1
            const int gv = g - 1, q = gv / kBPF, b0 = (gv - q * kBPF) * BQ, par = gv & 1;
            float r0 = PM[par][0], r1 = M1[par][0];
            for (int i = 0; i < BQ; i++) sy[q & 1][b0 + i] = biquad_step_x(sx[q & 1][b0 + i], r0, r1);
            if (__float_as_uint(r0) != __float_as_uint(PM[par][BQ]) || __float_as_uint(r1) != __float_as_uint(M1[par][1])) {
                // the fast step was wrong: block g (computed from its end state) is void, redo it from the true state
                m0 = r0; m1 = r1;
                if (g < G) {
                    const int q2 = g / kBPF, c0 = (g - q2 * kBPF) * BQ;
                    for (int i = 0; i < BQ; i++) sy[q2 & 1][c0 + i] = biquad_step_x(sx[q2 & 1][c0 + i], m0, m1);
                    FINAL[g & 1] = 1;
                }
            }
        }
        __syncthreads();                                           // replay results visible before anyone moves on
    }
    if (threadIdx.x == 0) { persist[n].biq[0] = m0; persist[n].biq[1] = m1; }
}

// Third form (knob bq_form=3; also an experiment that did not pay: 270 ms against 244 ms): the chain warp and the verifying
// warps no longer meet at two CTA-wide barriers per block.
// ncu on the second form (profiles/r02a: source page, per-instruction samples of the chain warp) put the chain loop at
// 35.4 cycles per sample and the per-block hand-over at ~640 cycles -- 6.7 cycles per sample at 96 samples per block: two
// __syncthreads round trips, the reconvergence of the divergent warp, instruction-cache misses on the jumps between the
// warps' code regions, the division of the block index, the exposed LDS latency of the block's first inputs.  Here
//  * warp 0 runs its own loop: block g, then a NON-blocking bar.arrive on FULL[g & 1]; the verifying warps run theirs:
//    bar.sync on FULL[g & 1], verify block g (hazard analysis, y stores, verdict flag), bar.arrive on DONE[g & 1].  The chain
//    warp reads the verdict on block g-1 after block g -- bar.sync on DONE[(g-1) & 1], which the verifiers reached one block
//    ago, so it never waits for them;
//  * the chain lane's inputs travel in three register stages (current iteration, next, next-but-one: 16 samples each),
//    fetched straight across block boundaries, so no LDS latency meets the chain whichever way ptxas orders the loads.
// A bad verdict on block g-1 (hazard, or a difference from the fast step): the chain warp first takes DONE[g & 1] as well
// (the verifiers may be busy with block g, which it has just computed from a state that may be wrong), replays block g-1 with
// the reference's double arithmetic from the recorded state and, only if the end state differs from what it carried on,
// recomputes block g the same way and voids the verifiers' verdict on it.  Frames leave for global memory two blocks behind
// the verifiers, when every block of the frame is final.  Blocks per frame >= 3 (BQ <= 160) keeps the frame buffers apart.
__device__ __forceinline__ void nb_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void nb_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <int BQ>
__global__ void __launch_bounds__(bq_threads(BQ)) k_rn_biquad_v3(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                                 RnPersist *persist, float *__restrict__ hp, size_t hp_stride, size_t N,
                                                                 int dbg) {
    // dbg (knob bq_dbg, measurement only, results then unverified): 1 = no hazard analysis, 2 = verdicts ignored
    constexpr int kBPF = kFrame / BQ, kThreads = bq_threads(BQ), kSub = 16, kIPB = BQ / kSub;
    static_assert(kFrame % BQ == 0 && BQ % kSub == 0 && kBPF >= 3, "blocks tile the frame, at least three per frame");
    constexpr int kFull = 1, kDone = 3;                    // named barriers: FULL[par] = 1 + par, DONE[par] = 3 + par
    const size_t n = blockIdx.x;
    if (n >= N) return;
    __shared__ __align__(16) float sx[3][kFrame];          // input scaled by 32767: frame q in sx[q % 3]
    __shared__ __align__(16) float sy[2][kFrame];
    __shared__ __align__(16) float PM[2][BQ + 4];          // m0 before every sample of the block; [BQ] = after the block
    __shared__ float M1[2][2];                             // m1 before / after the block
    __shared__ int FINAL[2], VERDICT[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *dst = hp + n * hp_stride + kPHist;
    auto src = [&](int q) { return in + (n * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame; };
    if (threadIdx.x < 2) { FINAL[threadIdx.x] = 0; VERDICT[threadIdx.x] = 0; }
    if (Q > 0) for (int i = threadIdx.x; i < kFrame; i += kThreads) sx[0][i] = src(0)[i] * 32767.0f;
    if (Q > 1) for (int i = threadIdx.x; i < kFrame; i += kThreads) sx[1][i] = src(1)[i] * 32767.0f;
    if (Q > 2) for (int i = threadIdx.x; i < kFrame; i += kThreads) sx[2][i] = src(2)[i] * 32767.0f;
    __syncthreads();
    const int G = Q * kBPF;

    if (warp == 0) {
        // ================================ the chain warp ================================
        float m0 = persist[n].biq[0], m1 = persist[n].biq[1];     // meaningful on lane 0
        float4 xa[kSub / 4], xb[kSub / 4], xc[kSub / 4];          // inputs of this iteration, the next, the one after
        bool primed = false;
        bool done_taken = false;                                  // DONE of the block about to be judged was taken early
        int q = 0, qb = 0, b0 = 0;                                // block g = frame q (buffer qb = q % 3), samples b0 .. b0 + BQ
        // Whole warp: the verdict on block gv (its verifiers have arrived on DONE[gv & 1], or are about to).
        auto judge = [&](int gv, int qv, int bv) {
            const int pv = gv & 1;
            if (!done_taken) nb_sync(kDone + pv, kThreads);
            done_taken = false;
            int bad = 0;
            if (lane == 0) bad = VERDICT[pv] && !(dbg & 2);
            bad = __shfl_sync(0xffffffffu, bad, 0);
            if (!bad) return;
            const bool has_next = gv + 1 < G;
            if (has_next) { nb_sync(kDone + (pv ^ 1), kThreads); done_taken = true; }   // nobody touches block gv+1 any more
            if (lane == 0) {
                VERDICT[pv] = 0;
                float r0 = PM[pv][0], r1 = M1[pv][0];
                for (int i = 0; i < BQ; i++) sy[qv & 1][bv + i] = biquad_step_x(sx[qv % 3][bv + i], r0, r1);
                if (__float_as_uint(r0) != __float_as_uint(PM[pv][BQ]) || __float_as_uint(r1) != __float_as_uint(M1[pv][1])) {
                    // the fast step was wrong: block gv+1 (computed from its end state) is void, redo it from the true state
                    m0 = r0; m1 = r1;
                    if (has_next) {
                        int q2 = qv, c0 = bv + BQ;
                        if (c0 == kFrame) { c0 = 0; q2++; }
                        for (int i = 0; i < BQ; i++) sy[q2 & 1][c0 + i] = biquad_step_x(sx[q2 % 3][c0 + i], m0, m1);
                        VERDICT[pv ^ 1] = 0;                      // whatever the verifiers found on the void data
                    }
                }
            }
            __syncwarp();
        };
        for (int g = 0; g < G; g++) {
            const int par = g & 1;
            if (lane == 0) {
                const float *xin = &sx[qb][b0];
                const float a0 = fabsf(m0), a1 = fabsf(m1);
                if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
                    float *yout = &sy[q & 1][b0];
                    for (int i = 0; i < BQ; i++) yout[i] = biquad_step_x(xin[i], m0, m1);     // denormal tail of a silence
                    FINAL[par] = 1;
                    primed = false;
                } else {
                    M1[par][0] = m1;
                    // position of the iteration after next, for the prefetch (runs on across block and frame boundaries)
                    int fq = q, fb = qb, fo = b0 + 2 * kSub;
                    if (fo >= kFrame) { fo -= kFrame; fq++; fb = fb == 2 ? 0 : fb + 1; }
                    if (!primed) {
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xa[k] = *reinterpret_cast<const float4 *>(xin + 4 * k);
                        int nq = q, nb = qb, no = b0 + kSub;
                        if (no >= kFrame) { no -= kFrame; nq++; nb = nb == 2 ? 0 : nb + 1; }
                        if (nq >= Q) { nb = qb; no = b0; }
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xb[k] = *reinterpret_cast<const float4 *>(&sx[nb][no + 4 * k]);
                        primed = true;
                    }
#pragma unroll 1
                    for (int it = 0; it < kIPB; it++) {
                        const float *xf = (fq < Q) ? &sx[fb][fo] : xin;         // past the end: a harmless re-fetch
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) xc[k] = *reinterpret_cast<const float4 *>(xf + 4 * k);
                        fo += kSub;
                        if (fo >= kFrame) { fo -= kFrame; fq++; fb = fb == 2 ? 0 : fb + 1; }
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) {
                            float4 pm;
                            pm.x = m0; biquad_chain_step(xa[k].x, m0, m1);
                            pm.y = m0; biquad_chain_step(xa[k].y, m0, m1);
                            pm.z = m0; biquad_chain_step(xa[k].z, m0, m1);
                            pm.w = m0; biquad_chain_step(xa[k].w, m0, m1);
                            *reinterpret_cast<float4 *>(&PM[par][it * kSub + 4 * k]) = pm;
                        }
#pragma unroll
                        for (int k = 0; k < kSub / 4; k++) { xa[k] = xb[k]; xb[k] = xc[k]; }
                    }
                    PM[par][BQ] = m0;
                    M1[par][1] = m1;
                    FINAL[par] = 0;
                }
            }
            __syncwarp();
            __threadfence_block();
            nb_arrive(kFull + par, kThreads);                      // block g is published; nobody is waited for
            if (g >= 1) {
                int qv = q, bv = b0 - BQ;
                if (bv < 0) { bv += kFrame; qv--; }
                judge(g - 1, qv, bv);      // (a replay leaves the register stages valid: they hold INPUTS of block g+1 onwards)
            }
            b0 += BQ;
            if (b0 == kFrame) { b0 = 0; q++; qb = qb == 2 ? 0 : qb + 1; }
        }
        if (G >= 1) {
            int qv = q, bv = b0 - BQ;
            if (bv < 0) { bv += kFrame; qv--; }
            judge(G - 1, qv, bv);
        }
        if (lane == 0) { persist[n].biq[0] = m0; persist[n].biq[1] = m1; }
    } else {
        // ================================ verifying warps (warp 1 also moves the frames) ================================
        const int v = (int)threadIdx.x - 32;                       // one sample of the block per thread
        float pf[kFrame / 32]; int pf_buf = -1;                    // warp 1: a frame in flight from global memory
        int q = 0, qb = 0, b0 = 0;
        for (int g = 0; g < G; g++) {
            const int par = g & 1;
            if (warp == 1 && pf_buf >= 0) {                        // the frame fetched one block ago lands in shared memory
#pragma unroll
                for (int k = 0; k < kFrame / 32; k++) sx[pf_buf][lane + 32 * k] = pf[k] * 32767.0f;
                pf_buf = -1;
            }
            nb_sync(kFull + par, kThreads);                        // block g is there
            bool hz = false;
            if (v < BQ && !FINAL[par]) {
                const float *xb = &sx[qb][b0];
                float a = PM[par][v], b;
                if (v == 0) b = M1[par][0];
                else { const float xp = xb[v - 1]; b = __fmaf_rn(-0.99600f, xp + PM[par][v - 1], xp); }
                const float xv = xb[v];
                float yv;
                if (dbg & 1) yv = xv + a;
                else {
                    yv = biquad_step_eft_x(xv, a, b, hz);
                    hz |= (__float_as_uint(a) != __float_as_uint(PM[par][v + 1])) |
                          (__float_as_uint(b) != __float_as_uint(__fmaf_rn(-0.99600f, yv, xv)));
                }
                sy[q & 1][b0 + v] = yv;                            // the reference's own float add: nothing to verify
            }
            if (__any_sync(0xffffffffu, hz) && lane == 0) VERDICT[par] = 1;
            if (warp == 1 && g >= 2) {                             // block g-2 is final: frame I/O at frame edges
                int bf = b0 - 2 * BQ, qf = q;
                if (bf < 0) { bf += kFrame; qf--; }
                if (bf == kFrame - BQ) {                           // it was the last block of frame qf
                    for (int i = lane; i < kFrame; i += 32) dst[(size_t)qf * kFrame + i] = sy[qf & 1][i];
                    // sx[qf % 3] is free: fetch frame qf + 3 (into registers now, into shared memory next block: the load
                    // latency never meets a barrier).  The chain reaches that frame more than a frame later, behind several
                    // DONE barriers that follow the write.
                    if (qf + 3 < Q) {
                        const float *s2 = src(qf + 3);
#pragma unroll
                        for (int k = 0; k < kFrame / 32; k++) pf[k] = s2[lane + 32 * k];
                        pf_buf = qf % 3;
                    }
                }
            }
            __threadfence_block();
            nb_arrive(kDone + par, kThreads);                      // done with block g (verdict, y, frame I/O)
            b0 += BQ;
            if (b0 == kFrame) { b0 = 0; q++; qb = qb == 2 ? 0 : qb + 1; }
        }
    }
    __syncthreads();                                               // every verdict is in, every replay written
    if (warp == 1) {                                               // the frames whose last block became final after the loops
        for (int qf = 0; qf < Q; qf++)
            if (qf * kBPF + kBPF - 1 + 2 >= G)
                for (int i = lane; i < kFrame; i += 32) dst[(size_t)qf * kFrame + i] = sy[qf & 1][i];
    }
}

// Fourth form (knob bq_form=4; the fifth form below, the default in the latency regime, is this one with a staging warp): the
// chain and its verification are SEPARATE KERNELS.  In the chain-per-CTA kernels above the lane's loop costs 41.6-42.6 cycles per sample
// whatever is switched off beside it; tools/ubench/chain_mix.cu runs the same step alone at 32-33.
//   k_rn_biquad_chain   one warp per state, lane 0 only: reads its PCM straight from global memory (16-byte loads one
//                       iteration ahead, L2 prefetch six frames ahead), runs biquad_chain_step, and publishes m0 before every
//                       sample (one 16-byte store per four samples) and m1 before every 8 samples.  No shared memory, no
//                       barrier, no hand-shake.  Iterations that start in a denormal tail take the reference's own double step.
//   k_rn_biquad_verify  frame-parallel, a warp per 32 samples: rebuilds each sample's m1 the way the lane formed it,
//                       re-evaluates the step with the full hazard analysis (biquad_step_eft_x), compares with the state
//                       the lane carried on, and stores y = x + m0 (the reference's own float add).  A warp that saw a
//                       hazard replays its 32 samples on one lane with the reference's double arithmetic from the recorded
//                       state; the replay ends in the recorded state of the next block unless the fast step was really
//                       wrong -- then the block's index goes to broken[state] (atomicMin).  0.6 ms per 20 x 60 s tile.
//   k_rn_biquad_fixup   a warp per state, returns at once unless broken[state] is set: everything from the first wrong
//                       block to the end of the tile is recomputed with the double step from the recorded (still true)
//                       state.  On the bench streams the fast step is never wrong outside denormal tails (0 of 86 M
//                       samples on the host), so this kernel is a launch and a load; knob bq_dbg=32 plants a wrong state
//                       to exercise it, bq_dbg=64 switches the repair off (tests/test_gpu_denoiser.py).
// The verdict no longer has to reach the lane while it runs, which is what tied the two together in one CTA.
// What the lane's loop taught (ncu source pages of four layouts, 20 streams x 60 s):
//  * ptxas puts all loads of the loop on ONE counting scoreboard: a first use that comes after younger loads were issued
//    waits for those too.  Loads re-issued quad by quad into the registers just consumed: 7 cycles per sample of
//    long-scoreboard stall (the full L2 latency every iteration); two register sets taking turns: the same, and plain loads
//    are sunk to their first use to save registers (volatile ones to the stores that follow them).  What works is the order
//    "scale the whole set (first use), THEN reload it", which the register dependence itself enforces.
//  * a taken branch costs the lone warp ~17 cycles (instruction fetch): the fast path closes the loop itself and the
//    denormal path sits behind it; the not-taken branch between head and body also keeps the head's loads at the head.
//  * what is left at the head while the chain idles: the scalings (packed, FMUL2: two samples per instruction), the 16-byte
//    loads (the LSU takes one every 4 cycles) and the pointer arithmetic.
// 39.4 cycles per sample (230.9 ms per 240 s of 20 streams, 24-sample iterations) against 42.0 for the first form; the loop
// body itself runs at 34.2.
constexpr int kBq4It = 8, kBq4ItPerFrame = kFrame / kBq4It, kBq4Blk = 32, kBq4VerWarps = 8;   // kBq4It: samples between two records of m1
constexpr int kBq4None = 0x7fffffff;
static_assert(kFrame % kBq4Blk == 0 && kBq4Blk % kBq4It == 0, "verify blocks tile the frame");
__device__ __forceinline__ bool bq4_tiny(float m0, float m1) {
    const float a0 = fabsf(m0), a1 = fabsf(m1);
    return (a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f);
}
// IT = samples per iteration of the lane's loop (16 ... 80).  At the head of an iteration the raw samples loaded one
// iteration ago are scaled into a second register set -- their first use, the only place the lane waits for memory -- and
// the loads of the NEXT iteration are issued at once into the registers just read: a whole iteration of chain time to land.  The order matters: ptxas puts all of a loop's loads on ONE counting scoreboard, so a first use that
// comes after younger loads were issued waits for those as well (ncu source page of two earlier layouts: 275 cycles of
// long-scoreboard stall per iteration, the full L2 latency, although the data had been requested 550 cycles earlier).
__device__ __forceinline__ float4 bq4_ld(const float *p) {
    float4 v;
    asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
// MONO: one channel per state, the state's frames are contiguous in its row (no pointer logic in the loop).  The denormal-tail
// path is marked unlikely so that the fast path falls through: a taken branch costs the lone warp ~17 cycles of instruction
// fetch, and the not-taken branch behind the head keeps the scaling and the loads in a basic block of their own -- at the head.
template <int IT, bool MONO>
__global__ void __launch_bounds__(32) k_rn_biquad_chain(const float *__restrict__ in, size_t in_stride, int C, int Q, size_t Tn,
                                                        RnPersist *persist, float *__restrict__ pm0, size_t pm0_stride,
                                                        float *__restrict__ pm1, size_t pm1_stride, int *__restrict__ broken,
                                                        size_t N, int dbg) {
    static_assert(kFrame % IT == 0 && IT % kBq4It == 0, "iterations tile the frame");
    constexpr int kQuads = IT / 4, kItPerFrame = kFrame / IT, kAheadFrames = 6, kAhead = kAheadFrames * kFrame;
    const size_t n = blockIdx.x;
    if (n >= N || threadIdx.x != 0) return;
    float m0 = persist[n].biq[0], m1 = persist[n].biq[1];
    broken[n] = kBq4None;
    float *mo = pm0 + n * pm0_stride, *m1o = pm1 + n * pm1_stride;
    const float *rowbase = in + n * (size_t)C * in_stride;
    if (Q <= 0) return;
    for (int c = 0; c < C; c++)                                     // the head of every channel row; the loop prefetches further on
        for (size_t a = 0; a < (size_t)kAhead && a < Tn * kFrame; a += 32)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(rowbase + (size_t)c * in_stride + a));
    float4 xr[kQuads];                                              // raw samples of the iteration about to run
#pragma unroll
    for (int k = 0; k < kQuads; k++) xr[k] = bq4_ld(rowbase + 4 * k);
    bool tiny = bq4_tiny(m0, m1);
    // the iteration being FETCHED: iteration nit of frame (time frame ntf, channel nc) of the state's sequence
    const float *pnext = rowbase;
    int nit = 0, nc = 0; size_t ntf = 0;
    const int H = Q * kItPerFrame;
    const int row_floats = (int)(Tn * kFrame);
    const int fault_h = ((dbg & 32) && n == 0) ? 40 * kItPerFrame - 1 : -1;    // test only: a wrong state after this iteration
    int h = 0;
    bool more;
    // scale the iteration's samples (their first use: the lane's only wait for memory), then fetch iteration h + 1 into the
    // registers just read (after the last iteration: any valid data)
    auto head = [&](float4 (&xs)[kQuads]) {
#pragma unroll
        for (int k = 0; k < kQuads; k++) {
            // two packed multiplies per quad (FMUL2: each half rounded like the scalar product): half the head's issue slots
            const float2 lo = __fmul2_rn(make_float2(xr[k].x, xr[k].y), make_float2(32767.0f, 32767.0f));
            const float2 hi = __fmul2_rn(make_float2(xr[k].z, xr[k].w), make_float2(32767.0f, 32767.0f));
            xs[k] = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
        more = h + 1 < H;
        const float *pl;
        if (MONO) {
            pnext += IT;
            pl = more ? pnext : rowbase;
            if ((h + 1) * IT + kAhead < row_floats) asm volatile("prefetch.global.L2 [%0];" ::"l"(pnext + kAhead));
        } else {
            nit++; pnext += IT;
            if (nit == kItPerFrame) {
                nit = 0; nc++;
                if (nc == C) { nc = 0; ntf++; }
                pnext = rowbase + (size_t)nc * in_stride + ntf * kFrame;
                if (ntf + kAheadFrames < Tn) {                      // a frame of this row, six frames ahead, into L2
                    const float *pf = pnext + kAhead;
#pragma unroll 1
                    for (int a = 0; a < kFrame; a += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + a));
                }
            }
            pl = more ? pnext : rowbase;
        }
#pragma unroll
        for (int k = 0; k < kQuads; k++) xr[k] = bq4_ld(pl + 4 * k);
    };
    auto foot = [&]() {
        if (h == fault_h) m0 = __uint_as_float(__float_as_uint(m0) ^ 0x1000u);
        mo += IT; m1o += IT / kBq4It; h++;
    };
#pragma unroll 1
    for (;;) {
        float4 xs[kQuads];
        head(xs);
        if (__builtin_expect(!tiny, 1)) {                           // (the branch keeps the head a basic block of its own: at the head)
#pragma unroll
            for (int k = 0; k < kQuads; k++) {
                if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                float4 pm;
                pm.x = m0; biquad_chain_step(xs[k].x, m0, m1);
                pm.y = m0; biquad_chain_step(xs[k].y, m0, m1);
                pm.z = m0; biquad_chain_step(xs[k].z, m0, m1);
                pm.w = m0; biquad_chain_step(xs[k].w, m0, m1);
                *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
                // decided four samples before the iteration ends (the state decays by 0.2 % per sample and the threshold
                // has five decades of margin): the branch behind the next head then never waits for the chain
                if (k == kQuads - 2) tiny = bq4_tiny(m0, m1);
            }
            foot();
            if (more) continue;                                     // the fast path closes the loop itself: one taken branch per iteration
            break;
        }
        {                                                           // denormal tail of a silence: the reference's own arithmetic
#pragma unroll
            for (int k = 0; k < kQuads; k++) {
                if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                float4 pm;
                pm.x = m0; biquad_step_x(xs[k].x, m0, m1);
                pm.y = m0; biquad_step_x(xs[k].y, m0, m1);
                pm.z = m0; biquad_step_x(xs[k].z, m0, m1);
                pm.w = m0; biquad_step_x(xs[k].w, m0, m1);
                *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
            }
            tiny = bq4_tiny(m0, m1);
        }
        foot();
        if (!more) break;
    }
    mo[0] = m0;                                                     // the end state closes both records
    m1o[0] = m1;
    persist[n].biq[0] = m0; persist[n].biq[1] = m1;
}

// Fifth form (knob bq_form=5): form 4 with a STAGING WARP.  Warp 1 copies the state's frames from global memory into a ring
// of four frames in shared memory, already scaled by 32767 (coalesced loads, 15 samples per lane and frame), up to three
// frames ahead of the chain; lane 0 of warp 0 reads its samples from the ring.  A shared-memory load lands in ~30 cycles, so
// the registers of a quad can be reloaded right after the quad has been consumed -- younger loads on the shared scoreboard
// are done long before the next first use -- and the loop has NO head: no scaling, no global loads, no pointer arithmetic on
// the chain's warp.  The two warps meet at no barrier: two counters in shared memory (frames staged / frames consumed), read
// once per frame.  Records, verification and fix-up are form 4's.  36.3 cycles per sample with 48-sample iterations (form 4:
// 39.4, form 1: 42.0); the body's asymptote is 35.7 -- what is left against the 28-cycle chain is ptxas' order of the lane's
// off-chain instructions.
constexpr int kBq5Ring = 4;
// a shared-memory load by its 32-bit shared address (formed ONCE: from a generic pointer the compiler re-derives the shared
// window with S2UR SR_CgaCtaId in every iteration, ~30 cycles the lane waits for)
__device__ __forceinline__ float4 bq5_lds(unsigned a) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
// One chain per CTA, written out on its own: the same source as the CH = 1 instance of the template below compiles to a
// different block layout there (the fast path loses its own back-edge: two taken branches per iteration, 40.3 instead of 36.8
// cycles per sample) -- the headline shape keeps the kernel whose code was measured.
template <int IT>
__global__ void __launch_bounds__(64) k_rn_biquad_chain_s1(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                          RnPersist *persist, float *__restrict__ pm0, size_t pm0_stride,
                                                          float *__restrict__ pm1, size_t pm1_stride, int *__restrict__ broken,
                                                          size_t N, int dbg) {
    static_assert(kFrame % IT == 0 && IT % kBq4It == 0, "iterations tile the frame");
    constexpr int kQuads = IT / 4, kItPerFrame = kFrame / IT, kRingFloats = kBq5Ring * kFrame;
    __shared__ __align__(16) float ring[kRingFloats];
    __shared__ int s_ready, s_done;                                 // frames staged / frames consumed so far
    const size_t n = blockIdx.x;
    if (n >= N) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { s_ready = 0; s_done = 0; broken[n] = kBq4None; }
    __syncthreads();
    volatile int *v_ready = &s_ready, *v_done = &s_done;
    if (warp == 1) {                                                // ---- the staging warp
        for (int q = 0; q < Q; q++) {
            const float *src = in + (n * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame;
            float v[kFrame / 32];
#pragma unroll
            for (int i = 0; i < kFrame / 32; i++) v[i] = __ldg(src + lane + 32 * i);
            while (q - *v_done >= kBq5Ring) __nanosleep(100);      // the slot is free once its frame has been consumed
            float *dst = ring + (q % kBq5Ring) * kFrame;
#pragma unroll
            for (int i = 0; i < kFrame / 32; i++) dst[lane + 32 * i] = v[i];
            __threadfence_block();
            __syncwarp();
            if (lane == 0) *v_ready = q + 1;
        }
        return;
    }
    if (lane != 0 || Q <= 0) return;
    // ---- the chain lane
    float m0 = persist[n].biq[0], m1 = persist[n].biq[1];
    float *mo = pm0 + n * pm0_stride, *m1o = pm1 + n * pm1_stride;
    while (*v_ready < 1) __nanosleep(20);
    __threadfence_block();
    auto scale = [](float4 v) {                                     // two packed multiplies per quad (each half rounded like the scalar product)
        const float2 lo = __fmul2_rn(make_float2(v.x, v.y), make_float2(32767.0f, 32767.0f));
        const float2 hi = __fmul2_rn(make_float2(v.z, v.w), make_float2(32767.0f, 32767.0f));
        return make_float4(lo.x, lo.y, hi.x, hi.y);
    };
    float4 xq[kQuads], xn[kQuads];                                  // scaled samples of the iteration about to run; raw samples of the next
#pragma unroll
    for (int k = 0; k < kQuads; k++) xq[k] = scale(*reinterpret_cast<const float4 *>(ring + 4 * k));
    bool tiny = bq4_tiny(m0, m1);
    unsigned ring_s;                                                // (opaque: the compiler would re-derive it inside the loop)
    asm volatile("mov.u32 %0, %1;" : "=r"(ring_s) : "r"((unsigned)__cvta_generic_to_shared(ring)));
    int off = 0;                                                    // ring offset of the iteration about to run
#pragma unroll 1
    for (int q = 0; q < Q; q++) {
        const int need = q + 2 < Q ? q + 2 : Q;                     // the frame's last iteration already reads the next frame
        while (*v_ready < need) __nanosleep(20);
        __threadfence_block();
        if ((dbg & 32) && n == 0 && q == 40) m0 = __uint_as_float(__float_as_uint(m0) ^ 0x1000u);   // test only: a wrong state
        int it = 0;
#pragma unroll 1
        for (;;) {
            off += IT;
            if (off == kRingFloats) off = 0;
            const unsigned nx = ring_s + 4u * (unsigned)off;        // after the very last iteration: stale data, never used
            it++;
#pragma unroll
            for (int k = 0; k < kQuads; k++) xn[k] = bq5_lds(nx + 16u * k);
            if (__builtin_expect(!tiny, 1)) {
#pragma unroll
                for (int k = 0; k < kQuads; k++) {
                    if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                    const float4 xs = xq[k];
                    float4 pm;
                    pm.x = m0; biquad_chain_step(xs.x, m0, m1);
                    pm.y = m0; biquad_chain_step(xs.y, m0, m1);
                    pm.z = m0; biquad_chain_step(xs.z, m0, m1);
                    pm.w = m0; biquad_chain_step(xs.w, m0, m1);
                    *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
                    if (k == kQuads - 2) tiny = bq4_tiny(m0, m1);   // four samples early: the branch never waits for the chain
                }
#pragma unroll
                for (int k = 0; k < kQuads; k++) xq[k] = scale(xn[k]);
                mo += IT; m1o += IT / kBq4It;
                if (it < kItPerFrame) continue;                     // the fast path closes the loop itself
                break;
            }
            {                                                       // denormal tail of a silence: the reference's own arithmetic
#pragma unroll
                for (int k = 0; k < kQuads; k++) {
                    if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                    const float4 xs = xq[k];
                    float4 pm;
                    pm.x = m0; biquad_step_x(xs.x, m0, m1);
                    pm.y = m0; biquad_step_x(xs.y, m0, m1);
                    pm.z = m0; biquad_step_x(xs.z, m0, m1);
                    pm.w = m0; biquad_step_x(xs.w, m0, m1);
                    *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
                }
                tiny = bq4_tiny(m0, m1);
#pragma unroll
                for (int k = 0; k < kQuads; k++) xq[k] = scale(xn[k]);
            }
            mo += IT; m1o += IT / kBq4It;
            if (it >= kItPerFrame) break;
        }
        *v_done = q + 1;
    }
    mo[0] = m0;                                                     // the end state closes both records
    m1o[0] = m1;
    persist[n].biq[0] = m0; persist[n].biq[1] = m1;
}

// CH = chains per CTA (1, 2 or 4): chain warps 0..CH-1 (one per SM sub-partition), staging warps CH..2CH-1, a ring and a pair
// of counters per chain.  One chain per SM is the fastest chain; two or four per SM leave more SMs to the other stages.
template <int IT, int CH>
__global__ void __launch_bounds__(64 * CH) k_rn_biquad_chain_s(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                               RnPersist *persist, float *__restrict__ pm0, size_t pm0_stride,
                                                               float *__restrict__ pm1, size_t pm1_stride, int *__restrict__ broken,
                                                               size_t N, int dbg) {
    static_assert(kFrame % IT == 0 && IT % kBq4It == 0, "iterations tile the frame");
    constexpr int kQuads = IT / 4, kItPerFrame = kFrame / IT, kRingFloats = kBq5Ring * kFrame;
    __shared__ __align__(16) float rings[CH][kRingFloats];
    __shared__ int s_ready[CH], s_done[CH];                         // frames staged / frames consumed so far, per chain
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // (the shuffle tells ptxas the value is warp-uniform)
    const int ch = warp < CH ? warp : warp - CH;
    const size_t n = (size_t)blockIdx.x * CH + ch;
    if (threadIdx.x < CH) {
        s_ready[threadIdx.x] = 0; s_done[threadIdx.x] = 0;
        if ((size_t)blockIdx.x * CH + threadIdx.x < N) broken[(size_t)blockIdx.x * CH + threadIdx.x] = kBq4None;
    }
    __syncthreads();
    if (n >= N) return;
    float *ring = rings[ch];
    volatile int *v_ready = &s_ready[ch], *v_done = &s_done[ch];
    if (warp >= CH) {                                               // ---- a staging warp
        for (int q = 0; q < Q; q++) {
            const float *src = in + (n * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame;
            float v[kFrame / 32];
#pragma unroll
            for (int i = 0; i < kFrame / 32; i++) v[i] = __ldg(src + lane + 32 * i);
            while (q - *v_done >= kBq5Ring) __nanosleep(100);      // the slot is free once its frame has been consumed
            float *dst = ring + (q % kBq5Ring) * kFrame;
#pragma unroll
            for (int i = 0; i < kFrame / 32; i++) dst[lane + 32 * i] = v[i];
            __threadfence_block();
            __syncwarp();
            if (lane == 0) *v_ready = q + 1;
        }
        return;
    }
    if (lane != 0 || Q <= 0) return;
    // ---- the chain lane
    float m0 = persist[n].biq[0], m1 = persist[n].biq[1];
    float *mo = pm0 + n * pm0_stride, *m1o = pm1 + n * pm1_stride;
    while (*v_ready < 1) __nanosleep(20);
    __threadfence_block();
    auto scale = [](float4 v) {                                     // two packed multiplies per quad (each half rounded like the scalar product)
        const float2 lo = __fmul2_rn(make_float2(v.x, v.y), make_float2(32767.0f, 32767.0f));
        const float2 hi = __fmul2_rn(make_float2(v.z, v.w), make_float2(32767.0f, 32767.0f));
        return make_float4(lo.x, lo.y, hi.x, hi.y);
    };
    float4 xq[kQuads], xn[kQuads];                                  // scaled samples of the iteration about to run; raw samples of the next
#pragma unroll
    for (int k = 0; k < kQuads; k++) xq[k] = scale(*reinterpret_cast<const float4 *>(ring + 4 * k));
    bool tiny = bq4_tiny(m0, m1);
    unsigned ring_s;                                                // (opaque: the compiler would re-derive it inside the loop)
    asm volatile("mov.u32 %0, %1;" : "=r"(ring_s) : "r"((unsigned)__cvta_generic_to_shared(ring)));
    int off = 0;                                                    // ring offset of the iteration about to run
#pragma unroll 1
    for (int q = 0; q < Q; q++) {
        const int need = q + 2 < Q ? q + 2 : Q;                     // the frame's last iteration already reads the next frame
        while (*v_ready < need) __nanosleep(20);
        __threadfence_block();
        if ((dbg & 32) && n == 0 && q == 40) m0 = __uint_as_float(__float_as_uint(m0) ^ 0x1000u);   // test only: a wrong state
        int it = 0;
#pragma unroll 1
        for (;;) {
            off += IT;
            if (off == kRingFloats) off = 0;
            const unsigned nx = ring_s + 4u * (unsigned)off;        // after the very last iteration: stale data, never used
            it++;
#pragma unroll
            for (int k = 0; k < kQuads; k++) xn[k] = bq5_lds(nx + 16u * k);
            if (__builtin_expect(!tiny, 1)) {
#pragma unroll
                for (int k = 0; k < kQuads; k++) {
                    if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                    const float4 xs = xq[k];
                    float4 pm;
                    pm.x = m0; biquad_chain_step(xs.x, m0, m1);
                    pm.y = m0; biquad_chain_step(xs.y, m0, m1);
                    pm.z = m0; biquad_chain_step(xs.z, m0, m1);
                    pm.w = m0; biquad_chain_step(xs.w, m0, m1);
                    *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
                    if (k == kQuads - 2) tiny = bq4_tiny(m0, m1);   // four samples early: the branch never waits for the chain
                }
#pragma unroll
                for (int k = 0; k < kQuads; k++) xq[k] = scale(xn[k]);
                mo += IT; m1o += IT / kBq4It;
                if (it < kItPerFrame) continue;                     // the fast path closes the loop itself
                break;
            }
            {                                                       // denormal tail of a silence: the reference's own arithmetic
#pragma unroll
                for (int k = 0; k < kQuads; k++) {
                    if (k % (kBq4It / 4) == 0) m1o[k / (kBq4It / 4)] = m1;
                    const float4 xs = xq[k];
                    float4 pm;
                    pm.x = m0; biquad_step_x(xs.x, m0, m1);
                    pm.y = m0; biquad_step_x(xs.y, m0, m1);
                    pm.z = m0; biquad_step_x(xs.z, m0, m1);
                    pm.w = m0; biquad_step_x(xs.w, m0, m1);
                    *reinterpret_cast<float4 *>(mo + 4 * k) = pm;
                }
                tiny = bq4_tiny(m0, m1);
#pragma unroll
                for (int k = 0; k < kQuads; k++) xq[k] = scale(xn[k]);
            }
            mo += IT; m1o += IT / kBq4It;
            if (it >= kItPerFrame) break;
        }
        *v_done = q + 1;
    }
    mo[0] = m0;                                                     // the end state closes both records
    m1o[0] = m1;
    persist[n].biq[0] = m0; persist[n].biq[1] = m1;
}

// sample t of state n's tile (frame q = t / 480 of the state's frame sequence (t0,c0),(t0,c1),...): scaled as the Denoiser does
__device__ __forceinline__ float bq4_x(const float *__restrict__ in, size_t in_stride, int C, size_t n, size_t t) {
    const size_t q = t / kFrame, i = t - q * kFrame;
    return in[(n * C + (q % C)) * in_stride + (q / C) * kFrame + i] * 32767.0f;
}
__global__ void __launch_bounds__(32 * kBq4VerWarps) k_rn_biquad_verify(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                                        const float *__restrict__ pm0, size_t pm0_stride,
                                                                        const float *__restrict__ pm1, size_t pm1_stride,
                                                                        int *__restrict__ broken, float *__restrict__ hp,
                                                                        size_t hp_stride, size_t N) {
    const size_t n = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const size_t vb = (size_t)blockIdx.x * kBq4VerWarps + (threadIdx.x >> 5);
    const size_t T = (size_t)Q * kFrame;
    if (n >= N || vb * kBq4Blk >= T) return;
    const size_t t = vb * kBq4Blk + lane;
    const float *M0 = pm0 + n * pm0_stride, *M1 = pm1 + n * pm1_stride;
    float *dst = hp + n * hp_stride + kPHist;
    const float xv = bq4_x(in, in_stride, C, n, t);
    const float pre0 = M0[t];
    const float xp = __shfl_up_sync(0xffffffffu, xv, 1), mp = __shfl_up_sync(0xffffffffu, pre0, 1);
    float a = pre0, b;
    if ((lane & (kBq4It - 1)) == 0) b = M1[t / kBq4It];
    else b = __fmaf_rn(-0.99600f, xp + mp, xp);                     // the lane's own m1' of the previous sample
    bool hz = false;
    const float yv = biquad_step_eft_x(xv, a, b, hz);
    hz |= __float_as_uint(a) != __float_as_uint(M0[t + 1]);
    const float b_lane = ((lane + 1) & (kBq4It - 1)) == 0 ? M1[(t + 1) / kBq4It] : __fmaf_rn(-0.99600f, yv, xv);
    hz |= __float_as_uint(b) != __float_as_uint(b_lane);
    if (!__any_sync(0xffffffffu, hz)) { dst[t] = yv; return; }
    if (lane == 0) {                                                // replay with the reference's double arithmetic
        float r0 = M0[t], r1 = M1[t / kBq4It];
        for (int i = 0; i < kBq4Blk; i++) dst[t + i] = biquad_step_x(bq4_x(in, in_stride, C, n, t + i), r0, r1);
        if (__float_as_uint(r0) != __float_as_uint(M0[t + kBq4Blk]) || __float_as_uint(r1) != __float_as_uint(M1[(t + kBq4Blk) / kBq4It]))
            atomicMin(&broken[n], (int)vb);                         // the fast step was wrong somewhere in this block
    }
}
__global__ void __launch_bounds__(32) k_rn_biquad_fixup(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                        const float *__restrict__ pm0, size_t pm0_stride,
                                                        const float *__restrict__ pm1, size_t pm1_stride,
                                                        const int *__restrict__ broken, RnPersist *persist,
                                                        float *__restrict__ hp, size_t hp_stride, size_t N) {
    const size_t n = blockIdx.x;
    if (n >= N || threadIdx.x != 0) return;
    const int kb = broken[n];
    if (kb == kBq4None) return;
    const size_t T = (size_t)Q * kFrame, t0 = (size_t)kb * kBq4Blk;
    float m0 = pm0[n * pm0_stride + t0], m1 = pm1[n * pm1_stride + t0 / kBq4It];   // every block before kb verified: this state is true
    float *dst = hp + n * hp_stride + kPHist;
    for (size_t t = t0; t < T; t++) dst[t] = biquad_step_x(bq4_x(in, in_stride, C, n, t), m0, m1);
    persist[n].biq[0] = m0; persist[n].biq[1] = m1;
}

// Throughput regime (many states): the same protocol with a state per LANE.  Warp 0 walks the chains of 32 states in
// lock-step (the fast step costs the warp the same instructions for 32 chains as for one -- with one chain per CTA the
// biquad was a quarter of all instructions the GPU issued at 592 streams), warps 1..7 verify the previous block (state =
// lane, the block's 32 samples dealt round-robin to the warps), store the block before that and fetch the next one.
// A state whose block raised a hazard replays it on its own lane with the reference's double arithmetic while the other
// lanes wait; the replay reproduces the fast result unless the fast step was really wrong, in which case that lane's
// block in flight is recomputed as well.  Blocks sit in shared memory as [state][36] floats (16-byte rows: the chain lane
// moves its 32 samples as eight 16-byte accesses, quarter-warp conflict-free); states before every sample as [sample][state].
constexpr int kBqlStates = 32, kBqlWarps = 8, kBqlStride = 36, kBqlVer = kBqlWarps - 1;
constexpr int kBqlBlock = 32, kBqlBlocksPerFrame = kFrame / kBqlBlock;
__global__ void __launch_bounds__(32 * kBqlWarps) k_rn_biquad_lanes(const float *__restrict__ in, size_t in_stride, int C, int Q,
                                                                    RnPersist *persist, float *__restrict__ hp, size_t hp_stride,
                                                                    size_t N) {
    __shared__ __align__(16) float sx[3][kBqlStates * kBqlStride];
    __shared__ __align__(16) float sy[3][kBqlStates * kBqlStride];
    __shared__ __align__(16) float2 P[2][(kBqlBlock + 1) * kBqlStates];
    __shared__ int HZ[kBqlStates];
    __shared__ int FINAL[2][kBqlStates];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t n0 = (size_t)blockIdx.x * kBqlStates;
    const int G = Q * kBqlBlocksPerFrame;
    const size_t n = n0 + lane;
    float m0 = 0.0f, m1 = 0.0f;                                    // meaningful on warp 0: lane l carries state n0 + l
    if (warp == 0 && n < N) { m0 = persist[n].biq[0]; m1 = persist[n].biq[1]; }
    if (warp == 0) { HZ[lane] = 0; FINAL[0][lane] = 0; FINAL[1][lane] = 0; }
    auto src = [&](size_t ns, int g) {
        const int q = g / kBqlBlocksPerFrame, b0 = (g - q * kBqlBlocksPerFrame) * kBqlBlock;
        return in + (ns * C + (q % C)) * in_stride + (size_t)(q / C) * kFrame + b0;
    };
    auto dst = [&](size_t ns, int g) {
        const int q = g / kBqlBlocksPerFrame, b0 = (g - q * kBqlBlocksPerFrame) * kBqlBlock;
        return hp + ns * hp_stride + kPHist + (size_t)q * kFrame + b0;
    };
    for (int s = warp; s < kBqlStates; s += kBqlWarps) sx[0][s * kBqlStride + lane] = (G > 0 && n0 + s < N) ? src(n0 + s, 0)[lane] : 0.0f;
    __syncthreads();
    for (int g = 0; g <= G + 1; g++) {
        if (warp == 0) {
            if (g < G) {                                           // ---- the chains: block g of 32 states
                const int par = g & 1;
                const float *xb = sx[g % 3] + lane * kBqlStride;
                float *yb = sy[g % 3] + lane * kBqlStride;
                float2 *Pb = P[par];
                float xr[kBqlBlock];
#pragma unroll
                for (int i = 0; i < kBqlBlock; i += 4) {
                    const float4 v = *reinterpret_cast<const float4 *>(xb + i);
                    xr[i] = v.x; xr[i + 1] = v.y; xr[i + 2] = v.z; xr[i + 3] = v.w;
                }
                const float a0 = fabsf(m0), a1 = fabsf(m1);
                const bool tiny = (a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f);
                if (__any_sync(0xffffffffu, tiny)) {
                    // A state in the denormal tail of a silence needs the reference's double arithmetic.  The warp would
                    // execute both paths one after the other, so ALL its states take the double path for this block: it is the
                    // reference's own arithmetic for every one of them, and one path (~85 cycles per sample) beats two.
#pragma unroll
                    for (int i = 0; i < kBqlBlock; i += 4) {
                        float4 yy;
                        yy.x = biquad_step(xr[i], m0, m1); yy.y = biquad_step(xr[i + 1], m0, m1);
                        yy.z = biquad_step(xr[i + 2], m0, m1); yy.w = biquad_step(xr[i + 3], m0, m1);
                        *reinterpret_cast<float4 *>(yb + i) = yy;
                    }
                    FINAL[par][lane] = 1;
                } else {
#pragma unroll
                    for (int i = 0; i < kBqlBlock; i += 4) {
                        float4 yy;
                        Pb[(i + 0) * kBqlStates + lane] = make_float2(m0, m1); yy.x = biquad_step_fast(xr[i], m0, m1);
                        Pb[(i + 1) * kBqlStates + lane] = make_float2(m0, m1); yy.y = biquad_step_fast(xr[i + 1], m0, m1);
                        Pb[(i + 2) * kBqlStates + lane] = make_float2(m0, m1); yy.z = biquad_step_fast(xr[i + 2], m0, m1);
                        Pb[(i + 3) * kBqlStates + lane] = make_float2(m0, m1); yy.w = biquad_step_fast(xr[i + 3], m0, m1);
                        *reinterpret_cast<float4 *>(yb + i) = yy;
                    }
                    Pb[kBqlBlock * kBqlStates + lane] = make_float2(m0, m1);
                    FINAL[par][lane] = 0;
                }
            }
        } else {
            const int vw = warp - 1;
            constexpr int kPer = (kBqlStates + kBqlVer - 1) / kBqlVer;
            float pre[kPer];                                        // block g+1 of this warp's states, in flight from global memory
            const bool fetch = g + 1 < G;
#pragma unroll
            for (int k = 0; k < kPer; k++) {
                const int s = vw + k * kBqlVer;
                pre[k] = (fetch && s < kBqlStates && n0 + s < N) ? src(n0 + s, g + 1)[lane] : 0.0f;
            }
            if (g >= 1 && g - 1 < G) {                              // ---- verification of block g-1: state = lane
                const int par = (g - 1) & 1;
                if (!FINAL[par][lane]) {
                    const float *xb = sx[(g - 1) % 3] + lane * kBqlStride, *yb = sy[(g - 1) % 3] + lane * kBqlStride;
                    for (int i = vw; i < kBqlBlock; i += kBqlVer) {
                        const float2 pre_s = P[par][i * kBqlStates + lane], post = P[par][(i + 1) * kBqlStates + lane];
                        float a = pre_s.x, b = pre_s.y;
                        bool hz = false;
                        const float yv = biquad_step_eft(xb[i], a, b, hz);
                        hz |= (__float_as_uint(a) != __float_as_uint(post.x)) | (__float_as_uint(b) != __float_as_uint(post.y)) |
                              (__float_as_uint(yv) != __float_as_uint(yb[i]));
                        if (hz) atomicOr(&HZ[lane], (int)(1u << i));      // which samples of this state's block: replayed one by one
                    }
                }
            }
            if (g >= 2) {                                           // ---- block g-2 is final: out it goes
                const float *yb = sy[(g - 2) % 3];
                for (int s = vw; s < kBqlStates; s += kBqlVer)
                    if (n0 + s < N) dst(n0 + s, g - 2)[lane] = yb[s * kBqlStride + lane];
            }
            if (fetch) {
#pragma unroll
                for (int k = 0; k < kPer; k++) {
                    const int s = vw + k * kBqlVer;
                    if (s < kBqlStates) sx[(g + 1) % 3][s * kBqlStride + lane] = pre[k];
                }
            }
        }
        __syncthreads();
        if (warp == 0) {
            unsigned hm = (g >= 1 && g - 1 < G) ? (unsigned)HZ[lane] : 0u;
            if (hm) {                                               // ---- replay of the flagged samples of block g-1, lane by lane
                // A hazard says the fast step could not be trusted at that sample; nearly always the double arithmetic
                // reproduces it and everything after it stands, so a flagged sample is redone on its own from its recorded
                // state (~100 cycles for the warp) and only a real difference voids the rest of the block and block g.
                const int par = (g - 1) & 1;
                const float *xb = sx[(g - 1) % 3] + lane * kBqlStride;
                float *yb = sy[(g - 1) % 3] + lane * kBqlStride;
                while (hm) {
                    const int v = __ffs((int)hm) - 1;
                    hm &= hm - 1;
                    float r0 = P[par][v * kBqlStates + lane].x, r1 = P[par][v * kBqlStates + lane].y;
                    yb[v] = biquad_step(xb[v], r0, r1);
                    const float2 post = P[par][(v + 1) * kBqlStates + lane];
                    if (__float_as_uint(r0) != __float_as_uint(post.x) || __float_as_uint(r1) != __float_as_uint(post.y)) {
                        for (int i = v + 1; i < kBqlBlock; i++) yb[i] = biquad_step(xb[i], r0, r1);
                        hm = 0;
                        const float2 end = P[par][kBqlBlock * kBqlStates + lane];
                        if (__float_as_uint(r0) != __float_as_uint(end.x) || __float_as_uint(r1) != __float_as_uint(end.y)) {
                            // the fast step was wrong: block g (computed from its end state) is void, redo it from the true state
                            m0 = r0; m1 = r1;
                            if (g < G) {
                                const float *xc = sx[g % 3] + lane * kBqlStride;
                                float *yc = sy[g % 3] + lane * kBqlStride;
                                for (int i = 0; i < kBqlBlock; i++) yc[i] = biquad_step(xc[i], m0, m1);
                                FINAL[g & 1][lane] = 1;
                            }
                        }
                    }
                }
            }
            HZ[lane] = 0;
        }
        __syncthreads();
    }
    if (warp == 0 && n < N) { persist[n].biq[0] = m0; persist[n].biq[1] = m1; }
}

constexpr int kWarpsPerCta = 4;     // 4 warps x <= 12 KB scratch: such a CTA still fits beside the RNN CTA (174 KB of weights) on an SM

__global__ void __launch_bounds__(32 * kWarpsPerCta) k_rn_analysis(Tables t, const float *__restrict__ hp, size_t hp_stride,
                                                                   size_t N, int Q, size_t Qcap, cpx *X, float *Ex,
                                                                   float *ceps, int *sil) {
    extern __shared__ float smem[];
    const int w = threadIdx.x >> 5;
    const size_t id = (size_t)blockIdx.x * kWarpsPerCta + w;
    if (id >= N * (size_t)Q) return;
    const size_t n = id / Q; const int q = (int)(id - n * Q);
    const size_t f = n * Qcap + q;
    analysis_warp(t, hp + n * hp_stride, q, X + f * kFreq, Ex + f * kNB, ceps + f * kNB, sil + f, smem + (size_t)w * kScrAna);
}

// F frames per CTA (F warps): the few-lane chains of the F frames are pooled (rn_core.cuh pitch_blk), so a larger F fills
// the pooled warps better -- 5F autocorrelation lanes, F scan lanes, 12F fine-search lanes.
template <int F>
__global__ void __launch_bounds__(32 * F) k_rn_pitch(Tables t, const float *__restrict__ hp, size_t hp_stride, size_t N,
                                                     int Q, size_t Qcap, PitchCand *cand) {
    extern __shared__ float smem[];
    (void)t;
    const float *xs[F];
    PitchCand *outs[F];
    const size_t id0 = (size_t)blockIdx.x * F, total = N * (size_t)Q;
#pragma unroll
    for (int f = 0; f < F; f++) {                           // a block past the end re-analyses its first frame, unwritten
        const bool ok = id0 + f < total;
        const size_t id = ok ? id0 + f : id0;
        const size_t n = id / Q; const int q = (int)(id - n * Q);
        xs[f] = hp + n * hp_stride + (size_t)q * kFrame;
        outs[f] = ok ? cand + n * Qcap + q : nullptr;
    }
    pitch_blk<F>(xs, outs, smem);
}

__global__ void k_rn_pitch_seq(const PitchCand *__restrict__ cand, size_t N, int Q, size_t Qcap, RnPersist *persist, int *pidx) {
    const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int lp = persist[n].last_period; float lg = persist[n].last_gain;
    for (int q = 0; q < Q; q++) {
        float g;
        const int idx = pitch_decide(cand[n * Qcap + q], lp, lg, &g);
        lp = idx; lg = g;                                     // denoise.c:334-335
        pidx[n * Qcap + q] = idx;
    }
    persist[n].last_period = lp; persist[n].last_gain = lg;
}

__global__ void __launch_bounds__(32 * kWarpsPerCta) k_rn_pfeat(Tables t, const float *__restrict__ hp, size_t hp_stride, size_t N,
                                                                int Q, size_t Qcap, const int *pidx, const cpx *X, const float *Ex,
                                                                cpx *P, float *Ep, float *Exp, float *featB) {
    extern __shared__ float smem[];
    const int w = threadIdx.x >> 5;
    const size_t id = (size_t)blockIdx.x * kWarpsPerCta + w;
    if (id >= N * (size_t)Q) return;
    const size_t n = id / Q; const int q = (int)(id - n * Q);
    const size_t f = n * Qcap + q;
    pfeat_warp(t, hp + n * hp_stride, q, pidx[f], X + f * kFreq, Ex + f * kNB, P + f * kFreq, Ep + f * kNB, Exp + f * kNB,
               featB + f * 8, smem + (size_t)w * kScrAna);
}

// ranks of non-silent frames: the cepstral ring advances only on them (denoise.c:358-369)
__global__ void k_rn_compact(const int *__restrict__ sil, size_t N, int Q, size_t Qcap, RnPersist *persist, long long *rank,
                             int *nzlist, int *nz_tile) {
    const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const long long base = persist[n].nz_total;
    int j = 0;
    for (int q = 0; q < Q; q++) {
        if (!sil[n * Qcap + q]) { rank[n * Qcap + q] = base + j; nzlist[n * Qcap + j] = q; j++; }
        else rank[n * Qcap + q] = -1;
    }
    nz_tile[n] = j;
    persist[n].nz_total = base + j;
}

__global__ void k_rn_compact_copy(const float *__restrict__ ceps, const int *__restrict__ nzlist, const int *__restrict__ nz_tile,
                                  size_t N, int Q, size_t Qcap, float *cepsc) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // one thread per (state, j, coefficient)
    const size_t per = (size_t)Q * kNB;
    if (id >= N * per) return;
    const size_t n = id / per; const size_t r = id - n * per;
    const int j = (int)(r / kNB), k = (int)(r - (size_t)j * kNB);
    if (j >= nz_tile[n]) return;
    const int q = nzlist[n * Qcap + j];
    cepsc[(n * (8 + Qcap) + 8 + j) * kNB + k] = ceps[(n * Qcap + q) * kNB + k];
}

__global__ void __launch_bounds__(32 * kWarpsPerCta) k_rn_cepsfeat(const float *__restrict__ cepsc, const int *__restrict__ nzlist,
                                                                   const long long *__restrict__ rank, const int *__restrict__ nz_tile,
                                                                   const float *featB, size_t N, int Q, size_t Qcap, float *feat) {
    __shared__ float smem[kWarpsPerCta * 80];
    const int w = threadIdx.x >> 5;
    const size_t id = (size_t)blockIdx.x * kWarpsPerCta + w;
    if (id >= N * (size_t)Q) return;
    const size_t n = id / Q; const int j = (int)(id - n * Q);
    if (j >= nz_tile[n]) return;
    const int q = nzlist[n * Qcap + j];
    const size_t f = n * Qcap + q;
    cepsfeat_warp(cepsc + (n * (8 + Qcap) + 8 + j) * kNB, rank[f], featB + f * 8, feat + f * kNFeat, smem + w * 80);
}

// One CTA = G independent groups of kRnnThreads threads; a group walks the non-silent frames of its R states in
// lock-step (round j = the j-th non-silent frame of each state; a state with fewer frames idles).  Weights (row-paired
// bf16, 174 KB) stay in shared memory, so there is one CTA per SM: G x R > 1 is what lets more than one state per SM make
// progress.  R states of a group share every weight fetch and unpack (fewer instructions per frame); G groups share only
// the weights and meet at no barrier after the start, so the narrow layers of one group (24..96 of its 288 threads busy)
// overlap the wide layers of another.  The frame index and the 42 features of the NEXT round are fetched while this
// round computes.
template <int R, int G, bool TOL = false>
__global__ void __launch_bounds__(kRnnThreads * G) k_rn_rnn(const unsigned *__restrict__ weights2, const float *__restrict__ bias,
                                                            const float *__restrict__ tansig, const float *__restrict__ feat,
                                                            const int *__restrict__ nzlist, const int *__restrict__ nz_tile,
                                                            RnPersist *persist, size_t Qcap, size_t N, float *gains, float *vad) {
    extern __shared__ __align__(16) unsigned char smraw[];
    unsigned *W2 = reinterpret_cast<unsigned *>(smraw);
    float *b = reinterpret_cast<float *>(smraw + ((kRnnWeights * 2 + 15) & ~15));
    float *ts = b + 560;
    const int grp = threadIdx.x / kRnnThreads, tid = threadIdx.x % kRnnThreads;
    float *scr = ts + 208 + (size_t)grp * R * kRv_end;
    const size_t n0 = ((size_t)blockIdx.x * G + grp) * R;
    for (int i = threadIdx.x; i < kRnnWeights / 2; i += kRnnThreads * G) W2[i] = weights2[i];
    for (int i = threadIdx.x; i < kRnnBiases; i += kRnnThreads * G) b[i] = bias[i];
    for (int i = threadIdx.x; i < 201; i += kRnnThreads * G) ts[i] = tansig[i];
    int nz[R], nzmax = 0;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const bool live = n0 + r < N;
        for (int i = tid; i < kRv_end; i += kRnnThreads) scr[r * kRv_end + i] = live ? persist[n0 + r].rnn_vec[i] : 0.f;
        nz[r] = live ? nz_tile[n0 + r] : 0;
        nzmax = nz[r] > nzmax ? nz[r] : nzmax;
    }
    // rounds j (current), j+1 (features in flight), j+2 (frame index in flight)
    int qa[R], qb[R];
    float fnext[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
        qa[r] = 0 < nz[r] ? nzlist[(n0 + r) * Qcap + 0] : 0;
        qb[r] = 1 < nz[r] ? nzlist[(n0 + r) * Qcap + 1] : 0;
        fnext[r] = (tid < kNFeat && 0 < nz[r]) ? feat[((n0 + r) * Qcap + qa[r]) * kNFeat + tid] : 0.f;
    }
    __syncthreads();                                        // the only CTA-wide barrier: weights are in place
    for (int j = 0; j < nzmax; j++) {
        unsigned act = 0;
        RnnOut<R> out;
#pragma unroll
        for (int r = 0; r < R; r++) {
            const bool on = j < nz[r];
            act |= on ? 1u << r : 0u;
            const size_t f = (n0 + r) * Qcap + qa[r];
            out.gains[r] = gains + f * kNB;
            out.vad[r] = vad + f;
            if (tid < kNFeat && on) scr[r * kRv_end + kRv_feat + tid] = fnext[r];
        }
        rnn_group_sync();
#pragma unroll
        for (int r = 0; r < R; r++) {                       // prefetch for the next two rounds
            if (tid < kNFeat && j + 1 < nz[r]) fnext[r] = feat[((n0 + r) * Qcap + qb[r]) * kNFeat + tid];
            qa[r] = qb[r];
            qb[r] = j + 2 < nz[r] ? nzlist[(n0 + r) * Qcap + j + 2] : 0;
        }
        rnn_frame_cta<R, TOL>(W2, b, ts, scr, act, out);
    }
#pragma unroll
    for (int r = 0; r < R; r++)
        if (n0 + r < N)
            for (int i = tid; i < kRv_end; i += kRnnThreads) persist[n0 + r].rnn_vec[i] = scr[r * kRv_end + i];
}

// lastg floor (denoise.c:478-482): one thread per (state, band); also zeroes vad/gains bookkeeping of silent frames
__global__ void k_rn_gain_scan(const float *__restrict__ gains, const int *__restrict__ sil, size_t N, int Q, size_t Qcap,
                               RnPersist *persist, float *gsm, float *vad) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= N * kNB) return;
    const size_t n = id / kNB; const int b = (int)(id - n * kNB);
    float lastg = persist[n].lastg[b];
    for (int q = 0; q < Q; q++) {
        const size_t f = n * Qcap + q;
        if (!sil[f]) {
            const float a = .6f * lastg;
            float g = gains[f * kNB + b];
            g = g > a ? g : a;
            lastg = g;
            gsm[f * kNB + b] = g;
        } else if (b == 0) vad[f] = 0.0f;                     // vad_prob stays 0 on silent frames (denoise.c:468)
    }
    persist[n].lastg[b] = lastg;
}

__global__ void __launch_bounds__(32 * kWarpsPerCta) k_rn_synth(Tables t, size_t N, int Q, size_t Qcap, const cpx *X, const cpx *P,
                                                                const float *Ex, const float *Ep, const float *Exp,
                                                                const float *gains, const float *gsm, const int *sil,
                                                                float *head, float *tail) {
    extern __shared__ float smem[];
    const int w = threadIdx.x >> 5;
    const size_t id = (size_t)blockIdx.x * kWarpsPerCta + w;
    if (id >= N * (size_t)Q) return;
    const size_t n = id / Q; const int q = (int)(id - n * Q);
    const size_t f = n * Qcap + q;
    synth_warp(t, X + f * kFreq, P + f * kFreq, Ex + f * kNB, Ep + f * kNB, Exp + f * kNB, gains + f * kNB, gsm + f * kNB, sil[f],
               head + f * kFrame, tail + (n * (1 + Qcap) + 1 + q) * kFrame, smem + (size_t)w * kScrSyn);
}

// overlap-add with the previous frame's tail, back to normalised PCM, channel rows (denoise.c:405-407, Denoiser.zig:88-94)
__global__ void k_rn_out(const float *__restrict__ head, const float *__restrict__ tail, size_t N, int C, int Q, size_t Qcap,
                         float *out, size_t out_stride, size_t out_off) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t per = (size_t)Q * kFrame;
    if (id >= N * per) return;
    const size_t n = id / per; const size_t r = id - n * per;
    const int q = (int)(r / kFrame), i = (int)(r - (size_t)q * kFrame);
    float v = head[(n * Qcap + q) * kFrame + i] + tail[(n * (1 + Qcap) + q) * kFrame + i];
    v *= (1.0f / 32767.0f);
    const int tt = q / C, c = q % C;
    out[(n * C + c) * out_stride + out_off + (size_t)tt * kFrame + i] = v;
}

// min over channels of the frame's VAD, start value 1 (VAD.zig:284-293)
__global__ void k_rn_vadmin(const float *__restrict__ vad, size_t N, int C, int Tn, size_t Qcap, float *vad_min, float *vad_all,
                            size_t vad_stride, size_t vad_off) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= N * (size_t)Tn) return;
    const size_t n = id / Tn; const int tt = (int)(id - n * Tn);
    float lo = 1.0f;
    for (int c = 0; c < C; c++) {
        const float v = vad[n * Qcap + (size_t)tt * C + c];
        if (v < lo) lo = v;
        if (vad_all) vad_all[(n * C + c) * vad_stride + vad_off + tt] = v;
    }
    if (vad_min) vad_min[n * vad_stride + vad_off + tt] = lo;
}

// Carries between tiles, one per stage so that the stages of different tiles can overlap.
// stage A: the last kPHist high-passed samples are the next tile's pitch/analysis history
__global__ void __launch_bounds__(256) k_rn_hp_save(const float *__restrict__ hp, size_t hp_stride, float *__restrict__ hist, int Q) {
    const size_t n = blockIdx.x;
    for (int i = threadIdx.x; i < kPHist; i += blockDim.x) hist[n * kPHist + i] = hp[n * hp_stride + (size_t)Q * kFrame + i];
}
__global__ void __launch_bounds__(256) k_rn_hp_load(float *__restrict__ hp, size_t hp_stride, const float *__restrict__ hist) {
    const size_t n = blockIdx.x;
    for (int i = threadIdx.x; i < kPHist; i += blockDim.x) hp[n * hp_stride + i] = hist[n * kPHist + i];
}
// stage B: the last 8 compacted cepstra move to the front (cepstral_mem, denoise.c:366-377)
__global__ void __launch_bounds__(256) k_rn_ceps_carry(float *cepsc, const int *__restrict__ nz_tile, size_t Qcap) {
    const size_t n = blockIdx.x;
    __shared__ float tmp[8 * kNB];
    const int nz = nz_tile[n];
    float *cc = cepsc + n * (8 + Qcap) * kNB;
    for (int i = threadIdx.x; i < 8 * kNB; i += blockDim.x) tmp[i] = cc[(size_t)nz * kNB + i];
    __syncthreads();
    for (int i = threadIdx.x; i < 8 * kNB; i += blockDim.x) cc[i] = tmp[i];
}
// stage D: the last frame's second half is the next tile's synthesis_mem (denoise.c:405-407)
__global__ void __launch_bounds__(256) k_rn_tail_carry(float *tail, int Q, size_t Qcap) {
    const size_t n = blockIdx.x;
    float *tl = tail + n * (1 + Qcap) * kFrame;
    for (int i = threadIdx.x; i < kFrame; i += blockDim.x) tl[i] = tl[(size_t)Q * kFrame + i];
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
#define FVAD_RNN_VARIANTS(X) X(1, 1) X(1, 2) X(1, 3) X(2, 1) X(2, 2) X(2, 3) X(4, 1) X(4, 2) X(4, 3)
// Dynamic shared memory a biquad CTA asks for without using it (latency regime): together with its static arrays it takes
// the SM's whole 227 KB, so that NO other CTA fits beside the chain (a co-resident frame-parallel CTA costs the chain ~13 %;
// with 200 KB the kernels that need no shared memory could still land there -- measured, filling the SM completely made no
// further difference: the slowdown then seen inside the wavefront was the instruction cache, see BQ / ROLL above).
static size_t biquad_smem_hog(size_t static_bytes) {
    const size_t per_block_max = 227 * 1024;
    return per_block_max > static_bytes ? per_block_max - static_bytes : 0;
}
static size_t g_bq_static[512];                              // static shared memory per BQ variant (cudaFuncGetAttributes)
static size_t g_bq2_static[512], g_bq3_static[512];
#define FVAD_BQ4_ITERS(X) X(16) X(24) X(32) X(48) X(80)
static int bq5_chains_per_sm(size_t N, size_t sms) { return 4 * N <= sms ? 1 : 2 * N <= sms ? 2 : 4; }   // latency regime: chains per reserved SM
static size_t bq5_static(int ch) { return (size_t)ch * (kBq5Ring * kFrame * 4 + 8) + 64; }   // upper bound of form 5's static shared memory
#define FVAD_BQ2_VARIANTS(X) X(48) X(96) X(160) X(240)
#define FVAD_BQ3_VARIANTS(X) X(48) X(96) X(160)
constexpr size_t kSmemAna = (size_t)kWarpsPerCta * kScrAna * sizeof(float);
constexpr size_t smem_pitch(int F) { return (size_t)F * kScrPitch * sizeof(float); }
#define FVAD_PITCH_VARIANTS(X) X(2) X(4) X(6) X(8)
constexpr size_t kSmemSyn = (size_t)kWarpsPerCta * kScrSyn * sizeof(float);
constexpr size_t smem_rnn(int states) { return ((kRnnWeights * 2 + 15) & ~15) + (560 + 208 + (size_t)states * kRv_end) * sizeof(float); }

template <class T> static int rn_alloc(RnnoiseBatch *r, T **p, size_t n) {
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T) + 256);
    if (e != cudaSuccess)
        return set_error(e == cudaErrorMemoryAllocation ? FVAD_EOUT_OF_MEMORY : FVAD_ECUDA, "cudaMalloc(%zu): %s", n * sizeof(T),
                         cudaGetErrorString(e));
    r->allocs.push_back(q);
    *p = (T *)q;
    return FVAD_OK;
}

void rnnoise_batch_destroy(RnnoiseBatch *r) {
    if (!r) return;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(r->device);
    for (void *p : r->allocs) cudaFree(p);
    if (prev >= 0) cudaSetDevice(prev);
    delete r;
}

int rnnoise_batch_reset(RnnoiseBatch *r, cudaStream_t stream) {
    FVAD_CUDA_TRY(cudaMemsetAsync(r->persist, 0, sizeof(RnPersist) * r->N, stream));
    FVAD_CUDA_TRY(cudaMemsetAsync(r->hp_hist, 0, sizeof(float) * r->N * kPHist, stream));
    FVAD_CUDA_TRY(cudaMemsetAsync(r->cepsc, 0, sizeof(float) * r->N * (8 + r->Qcap) * kNB, stream));
    FVAD_CUDA_TRY(cudaMemsetAsync(r->tail, 0, sizeof(float) * r->N * (1 + r->Qcap) * kFrame, stream));
    return FVAD_OK;
}

int rnnoise_batch_create(size_t n_states, size_t max_frames, int n_slots, int device, RnnoiseBatch **out) {
    *out = nullptr;
    if (!n_states || !max_frames || n_slots < 1) return set_error(FVAD_EINVALID_ARGUMENT, "rnnoise_batch_create: zero size");
    int prev = 0;
    FVAD_CUDA_TRY(cudaGetDevice(&prev));
    FVAD_CUDA_TRY(cudaSetDevice(device));
    RnnoiseBatch *r = new (std::nothrow) RnnoiseBatch();
    if (!r) return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed");
    r->device = device; r->N = n_states; r->Qcap = max_frames;
    int rc = rn_tables_get(device, &r->tab);
    const size_t NQ = r->N * r->Qcap;
    r->slots.resize(n_slots);
#define RN_A(p, n) if (rc == FVAD_OK) rc = rn_alloc(r, &p, (n))
    RN_A(r->persist, r->N);
    RN_A(r->hp_hist, r->N * kPHist);
    if (r->N < 1000) {                                            // the fourth biquad form's records (from 1000 states on: a state per lane)
        RN_A(r->bq_m0, r->N * r->bq_m0_stride()); RN_A(r->bq_m1, r->N * r->bq_m1_stride()); RN_A(r->bq_broken, r->N);
    }
    RN_A(r->cepsc, r->N * (8 + r->Qcap) * kNB);
    RN_A(r->head, NQ * kFrame); RN_A(r->tail, r->N * (1 + r->Qcap) * kFrame);
    for (RnSlot &s : r->slots) {
        RN_A(s.hp, r->N * r->hp_stride());
        RN_A(s.X, NQ * kFreq); RN_A(s.P, NQ * kFreq);
        RN_A(s.Ex, NQ * kNB); RN_A(s.Ep, NQ * kNB); RN_A(s.Exp, NQ * kNB); RN_A(s.ceps, NQ * kNB); RN_A(s.gains, NQ * kNB); RN_A(s.gsm, NQ * kNB);
        RN_A(s.featB, NQ * 8); RN_A(s.feat, NQ * kNFeat); RN_A(s.vad, NQ);
        RN_A(s.sil, NQ); RN_A(s.pidx, NQ); RN_A(s.nzlist, NQ); RN_A(s.rank, NQ); RN_A(s.cand, NQ); RN_A(s.nz_tile, r->N);
    }
#undef RN_A
    if (rc == FVAD_OK) {
        cudaError_t e = cudaFuncSetAttribute(k_rn_analysis, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemAna);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_pfeat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemAna);
#define FVAD_PITCH_ATTR(FF)                                                                                              \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_pitch<FF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pitch(FF));
        FVAD_PITCH_VARIANTS(FVAD_PITCH_ATTR)
#undef FVAD_PITCH_ATTR
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_synth, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemSyn);
#define FVAD_RNN_ATTR(RR, GG)                                                                                            \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_rnn<RR, GG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rnn(RR * GG));
        FVAD_RNN_VARIANTS(FVAD_RNN_ATTR)
#undef FVAD_RNN_ATTR
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_rnn<1, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rnn(1));
#define FVAD_BQ_ATTR(BB, RL)                                                                                             \
        if (e == cudaSuccess) {                                                                                          \
            cudaFuncAttributes fa{};                                                                                     \
            e = cudaFuncGetAttributes(&fa, k_rn_biquad<BB, RL>);                                                         \
            g_bq_static[BB] = fa.sharedSizeBytes;                                                                        \
            if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad<BB, RL>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                           (int)biquad_smem_hog(fa.sharedSizeBytes));                  \
        }
        FVAD_BQ_VARIANTS(FVAD_BQ_ATTR)
#undef FVAD_BQ_ATTR
#define FVAD_BQ2_ATTR(BB)                                                                                                \
        if (e == cudaSuccess) {                                                                                          \
            cudaFuncAttributes fa{};                                                                                     \
            e = cudaFuncGetAttributes(&fa, k_rn_biquad_v2<BB>);                                                          \
            g_bq2_static[BB] = fa.sharedSizeBytes;                                                                       \
            if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_v2<BB>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                           (int)biquad_smem_hog(fa.sharedSizeBytes));                  \
        }
        FVAD_BQ2_VARIANTS(FVAD_BQ2_ATTR)
#undef FVAD_BQ2_ATTR
#define FVAD_BQ3_ATTR(BB)                                                                                                \
        if (e == cudaSuccess) {                                                                                          \
            cudaFuncAttributes fa{};                                                                                     \
            e = cudaFuncGetAttributes(&fa, k_rn_biquad_v3<BB>);                                                          \
            g_bq3_static[BB] = fa.sharedSizeBytes;                                                                       \
            if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_v3<BB>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                           (int)biquad_smem_hog(fa.sharedSizeBytes));                  \
        }
        FVAD_BQ3_VARIANTS(FVAD_BQ3_ATTR)
#undef FVAD_BQ3_ATTR
#define FVAD_BQ4_ATTR(II)                                                                                                 \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_chain<II, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)biquad_smem_hog(0)); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_chain<II, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)biquad_smem_hog(0));
        FVAD_BQ4_ITERS(FVAD_BQ4_ATTR)
#undef FVAD_BQ4_ATTR
#define FVAD_BQ5_ATTR(II)                                                                                                 \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_chain_s1<II>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)biquad_smem_hog(bq5_static(1))); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_chain_s<II, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)biquad_smem_hog(bq5_static(2))); \
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_rn_biquad_chain_s<II, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)biquad_smem_hog(bq5_static(4)));
        FVAD_BQ4_ITERS(FVAD_BQ5_ATTR)
#undef FVAD_BQ5_ATTR
        if (e != cudaSuccess) rc = set_error(FVAD_ECUDA, "cudaFuncSetAttribute: %s (is this an sm_100 device?)", cudaGetErrorString(e));
    }
    if (rc == FVAD_OK) rc = rnnoise_batch_reset(r, 0);
    if (rc == FVAD_OK && cudaDeviceSynchronize() != cudaSuccess) rc = set_error(FVAD_ECUDA, "sync failed: %s", cudaGetErrorString(cudaGetLastError()));
    cudaSetDevice(prev);
    if (rc != FVAD_OK) { rnnoise_batch_destroy(r); return rc; }
    *out = r;
    return FVAD_OK;
}

int rnnoise_n_slots(const RnnoiseBatch *r) { return (int)r->slots.size(); }

static int rn_check(RnnoiseBatch *r, int slot, size_t Qs) {
    if (!r) return set_error(FVAD_ENOT_INITIALIZED, "denoiser not initialised");
    if (slot < 0 || slot >= (int)r->slots.size()) return set_error(FVAD_EINVALID_ARGUMENT, "bad slot %d", slot);
    if (Qs > r->Qcap) return set_error(FVAD_EINVALID_ARGUMENT, "rnnoise: %zu frames exceed capacity %zu", Qs, r->Qcap);
    return FVAD_OK;
}

// Stage A: high-pass biquad over the tile (sample-sequential per state).
int rnnoise_stage_a(RnnoiseBatch *r, int slot, const float *d_in, size_t in_stride, int C, size_t Tn, cudaStream_t st) {
    const size_t Qs = Tn * (size_t)C;
    FVAD_TRY(rn_check(r, slot, Qs));
    if (Qs == 0) return FVAD_OK;
    RnSlot &s = r->slots[slot];
    const size_t N = r->N;
    FVAD_LAUNCH(k_rn_hp_load, (unsigned)N, 256, 0, st, s.hp, r->hp_stride(), r->hp_hist);
    // Few chains (latency regime): biquad_smem_hog bytes of unused dynamic shared memory keep every other CTA off
    // the chain's SM -- the chain lane is issue-latency bound and loses ~13 % when frame-parallel kernels of other
    // stages share its scheduler.  Many chains (throughput regime): no reservation, so several chains and the
    // other stages' CTAs share an SM; reserving it there serialises the biquad against everything else.
    const bool lat = latency_regime(N, r->device);
    // A state per lane (k_rn_biquad_lanes: 8 x fewer instructions per sample, 32 chains per warp) from 700 states on.  Its
    // warp advances at the pace of its slowest lane (~230 cycles per sample with the synthetic streams' overlapping
    // silences: the whole warp takes the double path while any state is in a denormal tail), and stage A of tile t+1 follows
    // stage A of tile t, so it only pays where the tile is short enough for that chain not to set the pace.  Measured,
    // x real time, lanes / chain-per-CTA: 296 streams 70.1k / 103.2k, 592: 104.4k / 116.5k, 1184: 130.4k / 113.0k,
    // 2368: 134.2k / 113.8k.
    bool lanes = !lat && N >= 700;                                 // round 2, ms per 120 s, lanes / chain-per-CTA: 592 streams 616.6 / 565.3, 800: 709.2 / 788.3, 900: 812.9 / 886.7, 1184: 953.6 / 1143.6
    if (const char *e = debug_knob("biquad_lanes")) lanes = atoi(e) != 0;
    if (lanes) {
        FVAD_LAUNCH_NAMED("k_rn_biquad", k_rn_biquad_lanes, (unsigned)ceil_div(N, (size_t)kBqlStates), 32 * kBqlWarps, 0, st, d_in,
                          in_stride, C, (int)Qs, r->persist, s.hp, r->hp_stride(), N);
    } else {
        const size_t hog_on = lat ? 1 : 0;
        int BQ = 96;
        if (const char *e = debug_knob("bq_block")) BQ = atoi(e);
        bool roll = BQ > 60;
        if (const char *e = debug_knob("bq_roll")) roll = atoi(e) != 0;
        bool launched = false;
        int dbg = 0;
        if (const char *e = debug_knob("bq_dbg")) dbg = atoi(e);
        // experiment (knob bq_grid_min): pad the grid with CTAs that return at once -- B300_MICROARCH.md reports an issue
        // throttle on low-grid kernels that vanishes from 148 CTAs on
        size_t bq_grid = N;
        if (const char *e = debug_knob("bq_grid_min")) bq_grid = std::max<size_t>(N, (size_t)atoi(e));
        // knob bq_form: 5 = the chain lane fed by a staging warp, verification and fix-up in kernels of their own; 4 = the same
        // with the lane reading global memory itself; 1 = chain lane + verifying warps in one CTA; 2 = light chain lane; 3 =
        // named-barrier hand-over.  Form 5 where every chain has an SM to itself (latency regime; ms per 240 s of 20 streams:
        // form 5 212.9, form 4 230.9, form 1 246.6); with other stages' CTAs on the same SM a lone chain warp loses more than
        // form 1's CTA does (x real time, form 4 / form 1: 148 streams 66.8k / 72.2k, 296: 105.4k / 109.6k, 592: 114.5k / 118.3k).
        int form = lat ? 5 : 1;
        if (const char *e = debug_knob("bq_form")) form = atoi(e);
        if (form == 4 && ((reinterpret_cast<uintptr_t>(d_in) & 15) || (in_stride & 3))) form = 5;   // form 4's lane loads 16 bytes at a time
        if (form == 5 && !r->bq_m0) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_form=5: not available from 1000 states on");
        if (form == 5) {
            launched = true;
            const size_t T = Qs * kFrame;
            int it = 48;                                      // samples per loop iteration (ms per 240 s of 20 streams: 16: 220.5, 24: 251.6, 32: 218.2, 48: 212.9, 80: 253.5 -- 255 registers; the spread is ptxas' scheduling of the body)
            if (const char *e = debug_knob("bq_iter")) it = atoi(e);
            // chains per CTA = per reserved SM (knob bq_chains: 1, 2, 4).  A second chain on the SM costs the first nothing (its
            // warp has a sub-partition of its own), four cost 1.5 %; every SM not reserved goes to the other stages.  ms per
            // 240 s, 1 / 2 / 4 chains: 20 streams 289.6 / 289.4 / 293.9, 60: 290.1 / 281.3 / 288.7, 80: 316.7 / 290.3 / 290.8,
            // 96: 354.1 / 303.6 / 300.3, 112: 409.5 / 335.5 / 311.3.
            int chains = bq5_chains_per_sm(N, (size_t)device_sm_count(r->device));
            if (const char *e = debug_knob("bq_chains")) chains = atoi(e);
            bool go5 = false;
#define FVAD_BQ5_GO(II, CC)                                                                                              \
                FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad_chain_s<II, CC>), (unsigned)ceil_div(N, (size_t)CC), 64 * CC,     \
                                  hog_on * biquad_smem_hog(bq5_static(CC)), st, d_in, in_stride, C, (int)Qs, r->persist, r->bq_m0, \
                                  r->bq_m0_stride(), r->bq_m1, r->bq_m1_stride(), r->bq_broken, N, dbg)
#define FVAD_BQ5_PICK(II)                                                                                                 \
            if (!go5 && it == II) {                                                                                      \
                go5 = true;                                                                                              \
                if (chains == 4) FVAD_BQ5_GO(II, 4); else if (chains == 2) FVAD_BQ5_GO(II, 2);                               \
                else FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad_chain_s1<II>), (unsigned)N, 64, hog_on * biquad_smem_hog(bq5_static(1)), st, \
                                       d_in, in_stride, C, (int)Qs, r->persist, r->bq_m0, r->bq_m0_stride(), r->bq_m1, r->bq_m1_stride(),  \
                                       r->bq_broken, N, dbg);                                                            \
            }
            FVAD_BQ4_ITERS(FVAD_BQ5_PICK)
#undef FVAD_BQ5_GO
#undef FVAD_BQ5_PICK
            if (!go5) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_iter=%d: 16, 24, 32, 48 or 80", it);
            FVAD_LAUNCH(k_rn_biquad_verify, dim3((unsigned)ceil_div(T / kBq4Blk, (size_t)kBq4VerWarps), (unsigned)N), 32 * kBq4VerWarps, 0, st,
                        d_in, in_stride, C, (int)Qs, r->bq_m0, r->bq_m0_stride(), r->bq_m1, r->bq_m1_stride(), r->bq_broken, s.hp,
                        r->hp_stride(), N);
            if (!(dbg & 64))
            FVAD_LAUNCH(k_rn_biquad_fixup, (unsigned)N, 32, 0, st, d_in, in_stride, C, (int)Qs, r->bq_m0, r->bq_m0_stride(), r->bq_m1,
                        r->bq_m1_stride(), r->bq_broken, r->persist, s.hp, r->hp_stride(), N);
        }
        if (form == 4 && !r->bq_m0) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_form=4: not available from 1000 states on");
        if (form == 4) {
            launched = true;
            const size_t T = Qs * kFrame;
            int it = 24;                                      // knob bq_iter: samples per iteration of the lane's loop (measured, ms per 240 s of 20 streams: 16: 241.7, 24: 230.9, 32: 243.0, 48: 230.5; the first form: 246.6 -- beyond the head's amortisation the differences are ptxas' scheduling of the body)
            if (const char *e = debug_knob("bq_iter")) it = atoi(e);
#define FVAD_BQ4_GO(II, MM)                                                                                              \
            FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad_chain<II, MM>), (unsigned)N, 32, hog_on * biquad_smem_hog(0), st, d_in, in_stride, C, \
                              (int)Qs, Tn, r->persist, r->bq_m0, r->bq_m0_stride(), r->bq_m1, r->bq_m1_stride(), r->bq_broken, N, dbg)
            bool go4 = false;
#define FVAD_BQ4_PICK(II)                                                                                                 \
            if (!go4 && it == II) { go4 = true; if (C == 1) FVAD_BQ4_GO(II, true); else FVAD_BQ4_GO(II, false); }
            FVAD_BQ4_ITERS(FVAD_BQ4_PICK)
#undef FVAD_BQ4_PICK
            if (!go4) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_iter=%d: 16, 24, 32, 48 or 80", it);
#undef FVAD_BQ4_GO
            FVAD_LAUNCH(k_rn_biquad_verify, dim3((unsigned)ceil_div(T / kBq4Blk, (size_t)kBq4VerWarps), (unsigned)N), 32 * kBq4VerWarps, 0, st,
                        d_in, in_stride, C, (int)Qs, r->bq_m0, r->bq_m0_stride(), r->bq_m1, r->bq_m1_stride(), r->bq_broken, s.hp,
                        r->hp_stride(), N);
            if (!(dbg & 64))                                   // (test only: bq_dbg=64 leaves a planted fault unrepaired)
            FVAD_LAUNCH(k_rn_biquad_fixup, (unsigned)N, 32, 0, st, d_in, in_stride, C, (int)Qs, r->bq_m0, r->bq_m0_stride(), r->bq_m1,
                        r->bq_m1_stride(), r->bq_broken, r->persist, s.hp, r->hp_stride(), N);
        }
        const bool v2 = form == 2;
#define FVAD_BQ3_GO(BB)                                                                                                  \
        if (form == 3 && !launched && BQ == BB) {                                                                        \
            launched = true;                                                                                             \
            FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad_v3<BB>), (unsigned)N, bq_threads(BB), hog_on * biquad_smem_hog(g_bq3_static[BB]), st, d_in, in_stride, C, (int)Qs,   \
                              r->persist, s.hp, r->hp_stride(), N, dbg);                                                 \
        }
        FVAD_BQ3_VARIANTS(FVAD_BQ3_GO)
#undef FVAD_BQ3_GO
        if (form == 3 && !launched) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_block=%d: no such variant of the third form", BQ);
#define FVAD_BQ2_GO(BB)                                                                                                  \
        if (v2 && !launched && BQ == BB) {                                                                               \
            launched = true;                                                                                             \
            FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad_v2<BB>), (unsigned)bq_grid, bq_threads(BB), hog_on * biquad_smem_hog(g_bq2_static[BB]), st, d_in, in_stride, C, (int)Qs,   \
                              r->persist, s.hp, r->hp_stride(), N, dbg);                                                    \
        }
        FVAD_BQ2_VARIANTS(FVAD_BQ2_GO)
#undef FVAD_BQ2_GO
#define FVAD_BQ_GO(BB, RL)                                                                                               \
        if (!launched && BQ == BB && roll == RL) {                                                                       \
            launched = true;                                                                                             \
            FVAD_LAUNCH_NAMED("k_rn_biquad", (k_rn_biquad<BB, RL>), (unsigned)bq_grid, bq_threads(BB), hog_on * biquad_smem_hog(g_bq_static[BB]), st, d_in, in_stride, C, (int)Qs,   \
                              r->persist, s.hp, r->hp_stride(), N);                                                      \
        }
        FVAD_BQ_VARIANTS(FVAD_BQ_GO)
#undef FVAD_BQ_GO
        if (!launched) return set_error(FVAD_EINVALID_ARGUMENT, "knob bq_block=%d: no such biquad variant", BQ);
    }
    FVAD_LAUNCH(k_rn_hp_save, (unsigned)N, 256, 0, st, s.hp, r->hp_stride(), r->hp_hist, (int)Qs);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

// Frames per pitch CTA.  Measured at 592 streams: F = 4 95.2k, 6 93.9k, 8 92.4k x real time -- the fuller pooled warps of
// a larger F do not pay for the longer barrier waits and the coarser shared-memory granularity (knob pitch_f overrides).
static int pitch_frames_per_cta(size_t, int) { return 4; }

// Stage B1: frame-parallel analysis and pitch correlations (no carry between tiles).
int rnnoise_stage_b1(RnnoiseBatch *r, int slot, size_t Qs, cudaStream_t st) {
    FVAD_TRY(rn_check(r, slot, Qs));
    if (Qs == 0) return FVAD_OK;
    RnSlot &s = r->slots[slot];
    const int Q = (int)Qs;
    const size_t N = r->N, Qcap = r->Qcap, NQ = N * Qs;
    const Tables &t = r->tab.t;
    const unsigned warp_grid = (unsigned)ceil_div(NQ, kWarpsPerCta);
    FVAD_LAUNCH(k_rn_analysis, warp_grid, 32 * kWarpsPerCta, kSmemAna, st, t, s.hp, r->hp_stride(), N, Q, Qcap, s.X, s.Ex, s.ceps, s.sil);
    {
        int F = pitch_frames_per_cta(NQ, r->device);
        if (const char *e = debug_knob("pitch_f")) { const int v = atoi(e); if (v == 2 || v == 4 || v == 6 || v == 8) F = v; }
        bool launched = false;
#define FVAD_PITCH_GO(FF)                                                                                                \
        if (!launched && F == FF) {                                                                                      \
            launched = true;                                                                                             \
            FVAD_LAUNCH_NAMED("k_rn_pitch", k_rn_pitch<FF>, (unsigned)ceil_div(NQ, (size_t)FF), 32 * FF, smem_pitch(FF), st, t, s.hp, \
                              r->hp_stride(), N, Q, Qcap, s.cand);                                                       \
        }
        FVAD_PITCH_VARIANTS(FVAD_PITCH_GO)
#undef FVAD_PITCH_GO
    }
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

// Stage B2: the pitch decision scan (frame-sequential, tiny) and what depends on it: pitch-lagged spectrum and features,
// the non-silent frame ranks and the cepstral-ring features.  Carries: last_period/last_gain, nz_total, cepstral_mem.
int rnnoise_stage_b2(RnnoiseBatch *r, int slot, size_t Qs, cudaStream_t st) {
    FVAD_TRY(rn_check(r, slot, Qs));
    if (Qs == 0) return FVAD_OK;
    RnSlot &s = r->slots[slot];
    const int Q = (int)Qs;
    const size_t N = r->N, Qcap = r->Qcap, NQ = N * Qs;
    const Tables &t = r->tab.t;
    const unsigned warp_grid = (unsigned)ceil_div(NQ, kWarpsPerCta);
    FVAD_LAUNCH(k_rn_pitch_seq, (unsigned)ceil_div(N, 32), 32, 0, st, s.cand, N, Q, Qcap, r->persist, s.pidx);
    FVAD_LAUNCH(k_rn_pfeat, warp_grid, 32 * kWarpsPerCta, kSmemAna, st, t, s.hp, r->hp_stride(), N, Q, Qcap, s.pidx, s.X, s.Ex, s.P,
                s.Ep, s.Exp, s.featB);
    FVAD_LAUNCH(k_rn_compact, (unsigned)ceil_div(N, 32), 32, 0, st, s.sil, N, Q, Qcap, r->persist, s.rank, s.nzlist, s.nz_tile);
    FVAD_LAUNCH(k_rn_compact_copy, (unsigned)ceil_div(NQ * kNB, 256), 256, 0, st, s.ceps, s.nzlist, s.nz_tile, N, Q, Qcap, r->cepsc);
    FVAD_LAUNCH(k_rn_cepsfeat, warp_grid, 32 * kWarpsPerCta, 0, st, r->cepsc, s.nzlist, s.rank, s.nz_tile, s.featB, N, Q, Qcap, s.feat);
    FVAD_LAUNCH(k_rn_ceps_carry, (unsigned)N, 256, 0, st, r->cepsc, s.nz_tile, Qcap);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

int rnnoise_stage_b(RnnoiseBatch *r, int slot, size_t Qs, cudaStream_t st) {
    FVAD_TRY(rnnoise_stage_b1(r, slot, Qs, st));
    return rnnoise_stage_b2(r, slot, Qs, st);
}

// Stage C: the RNN, frame-sequential per state.
int rnnoise_stage_c(RnnoiseBatch *r, int slot, size_t Qs, cudaStream_t st) {
    int tol_mode = 0;                                           // experiment (knob rnn_tolerance_mode; 0 = the product path): rn_core.cuh
    if (const char *e = debug_knob("rnn_tolerance_mode")) tol_mode = atoi(e);
    if (tol_mode) {
        static int set_mode[64] = {};
        if (r && r->device >= 0 && r->device < 64 && set_mode[r->device] != tol_mode) {
            FVAD_CUDA_TRY(cudaMemcpyToSymbolAsync(rn::c_rnn_tolerance_mode, &tol_mode, sizeof tol_mode, 0, cudaMemcpyHostToDevice, st));
            set_mode[r->device] = tol_mode;
        }
    }
    FVAD_TRY(rn_check(r, slot, Qs));
    if (Qs == 0) return FVAD_OK;
    RnSlot &s = r->slots[slot];
    // Latency regime (a state per SM at most): the shortest chain wins, R = G = 1.  Throughput regime: R states share a
    // group's weight fetches, G groups fill each other's narrow layers, and packing six states into a CTA (R, G = 2, 3)
    // leaves the other SMs wholly to the frame-parallel kernels (an RNN CTA's 174 KB of weights keep almost everything
    // else off its SM) -- measured, x real time: 148 streams (1,1) 81.3k, (1,3) 82.7k, (2,3) 87.8k; 296 streams (1,2)
    // 102.2k, (2,1) 100.2k, (2,3) 103.9k.  knob rnn_r / knob rnn_g override (tests exercise every variant).
    const size_t sms = (size_t)device_sm_count(r->device);
    const size_t per_sm = ceil_div(r->N, sms);
    int R = 1, G = 1;
    if (!latency_regime(r->N, r->device)) {
        if (per_sm <= 6) { R = 2; G = 3; }
        else { R = 4; G = 3; }                                  // 1184 streams, ms per 120 s: (4,3) 953.7, (2,3) 962.2, (4,2) 987.8, (2,2) 1008.5, (1,3) 1008.6, (4,1) 1140.3
    } else if (5 * r->N > 3 * sms) {
        // Upper latency regime: the biquad chains hold reserved SMs (four chains each up here), and an RNN CTA takes an SM of
        // the rest for its weights.  Two states per CTA (independent groups) from 0.6 states per SM on; ms per 240 s, G = 1 /
        // 2 / 3: 80 streams 285.9 / 290.7 / 292.2, 112: 334.0 / 312.4 / 312.9, 128: 369.3 / 328.1 / 320.8, 148: 411.5 / 370.2 /
        // 363.0.
        G = 2;
    }
    if (const char *e = debug_knob("rnn_r")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4) R = v; }
    if (const char *e = debug_knob("rnn_g")) { const int v = atoi(e); if (v >= 1 && v <= 3) G = v; }
    if (tol_mode) { R = 1; G = 1; }                             // the experiment's kernel exists in one shape
    const unsigned grid = (unsigned)ceil_div(r->N, (size_t)R * G);
    bool launched = false;
    if (tol_mode) {
        launched = true;
        FVAD_LAUNCH_NAMED("k_rn_rnn", (k_rn_rnn<1, 1, true>), grid, kRnnThreads, smem_rnn(1), st, r->tab.weights2, r->tab.bias, r->tab.t.tansig,
                          s.feat, s.nzlist, s.nz_tile, r->persist, r->Qcap, r->N, s.gains, s.vad);
    }
#define FVAD_RNN_GO(RR, GG)                                                                                              \
    if (!launched && R == RR && G == GG) {                                                                               \
        launched = true;                                                                                                 \
        FVAD_LAUNCH_NAMED("k_rn_rnn", (k_rn_rnn<RR, GG>), grid, kRnnThreads * GG, smem_rnn(RR * GG), st, r->tab.weights2,  \
                          r->tab.bias, r->tab.t.tansig, s.feat, s.nzlist, s.nz_tile, r->persist, r->Qcap, r->N, s.gains, s.vad); \
    }
    FVAD_RNN_VARIANTS(FVAD_RNN_GO)
#undef FVAD_RNN_GO
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

// Stage D: gain smoothing scan, synthesis, overlap-add to normalised PCM, per-frame VAD minimum.
int rnnoise_stage_d(RnnoiseBatch *r, int slot, int C, size_t Tn, float *d_out, size_t out_stride, size_t out_off, float *d_vad_min,
                    float *d_vad_all, size_t vad_stride, size_t vad_off, cudaStream_t st) {
    const size_t Qs = Tn * (size_t)C;
    FVAD_TRY(rn_check(r, slot, Qs));
    if (Qs == 0) return FVAD_OK;
    RnSlot &s = r->slots[slot];
    const int Q = (int)Qs;
    const size_t N = r->N, Qcap = r->Qcap, NQ = N * Qs;
    const unsigned warp_grid = (unsigned)ceil_div(NQ, kWarpsPerCta);
    FVAD_LAUNCH(k_rn_gain_scan, (unsigned)ceil_div(N * kNB, 64), 64, 0, st, s.gains, s.sil, N, Q, Qcap, r->persist, s.gsm, s.vad);
    FVAD_LAUNCH(k_rn_synth, warp_grid, 32 * kWarpsPerCta, kSmemSyn, st, r->tab.t, N, Q, Qcap, s.X, s.P, s.Ex, s.Ep, s.Exp, s.gains, s.gsm,
                s.sil, r->head, r->tail);
    FVAD_LAUNCH(k_rn_out, (unsigned)ceil_div(NQ * kFrame, 256), 256, 0, st, r->head, r->tail, N, C, Q, Qcap, d_out, out_stride, out_off);
    if (d_vad_min || d_vad_all)
        FVAD_LAUNCH(k_rn_vadmin, (unsigned)ceil_div(N * Tn, 128), 128, 0, st, s.vad, N, C, (int)Tn, Qcap, d_vad_min, d_vad_all, vad_stride, vad_off);
    FVAD_LAUNCH(k_rn_tail_carry, (unsigned)N, 256, 0, st, r->tail, Q, Qcap);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

int rnnoise_batch_process(RnnoiseBatch *r, const float *d_in, size_t in_stride, int C, size_t Tn, float *d_out,
                          size_t out_stride, size_t out_off, float *d_vad_min, float *d_vad_all, size_t vad_stride,
                          size_t vad_off, cudaStream_t st) {
    FVAD_TRY(rnnoise_stage_a(r, 0, d_in, in_stride, C, Tn, st));
    FVAD_TRY(rnnoise_stage_b(r, 0, Tn * (size_t)C, st));
    FVAD_TRY(rnnoise_stage_c(r, 0, Tn * (size_t)C, st));
    return rnnoise_stage_d(r, 0, C, Tn, d_out, out_stride, out_off, d_vad_min, d_vad_all, vad_stride, vad_off, st);
}

}  // namespace fvad

// ------------------------------------------------------------------------------------------
// test aid: the device's Ly = log10(1e-2 + Ex) (rn_core.cuh log10_ex) on an array
// ------------------------------------------------------------------------------------------
namespace fvad {
__global__ void k_debug_log10(const float *__restrict__ in, float *__restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = rn::log10_ex(in[i]);
}
}  // namespace fvad
extern "C" int fvad_debug_log10(const float *ex, float *out, size_t n) {
    if ((!ex || !out) && n) return fvad::set_error(FVAD_EINVALID_ARGUMENT, "fvad_debug_log10: null argument");
    if (!n) return FVAD_OK;
    float *d_in = nullptr, *d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_in, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(d_in, ex, n * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        FVAD_LAUNCH(fvad::k_debug_log10, 1184, 256, 0, 0, d_in, d_out, n);
        e = cudaMemcpy(out, d_out, n * sizeof(float), cudaMemcpyDeviceToHost);
    }
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return fvad::set_error(FVAD_ECUDA, "fvad_debug_log10: %s (no CPU fallback exists)", cudaGetErrorString(e));
    return FVAD_OK;
}

// ------------------------------------------------------------------------------------------
// C ABI: Denoiser.zig drop-ins
// ------------------------------------------------------------------------------------------
using namespace fvad;

struct fvad_denoiser_batch {
    RnnoiseBatch *rn = nullptr;
    size_t N = 0, Qcap = 0;
    int device = 0;
    float *d_in = nullptr, *d_out = nullptr, *d_vad = nullptr;
};
struct fvad_denoiser { fvad_denoiser_batch *b; float in[480]; };

extern "C" size_t fvad_denoiser_frame_size(void) { return 480; }       // Denoiser.zig:68-70

extern "C" void fvad_denoiser_batch_destroy(fvad_denoiser_batch *d) {
    if (!d) return;
    int prev = -1; cudaGetDevice(&prev); cudaSetDevice(d->device);
    rnnoise_batch_destroy(d->rn);
    cudaFree(d->d_in); cudaFree(d->d_out); cudaFree(d->d_vad);
    if (prev >= 0) cudaSetDevice(prev);
    delete d;
}

extern "C" int fvad_denoiser_batch_create(size_t n_streams, int device, fvad_denoiser_batch **out) {
    if (!out || !n_streams) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_denoiser_batch_create: bad argument");
    *out = nullptr;
    if (device < 0) {
        cudaError_t e = cudaGetDevice(&device);
        if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    }
    fvad_denoiser_batch *d = new (std::nothrow) fvad_denoiser_batch();
    if (!d) return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed");
    d->N = n_streams; d->device = device;
    d->Qcap = std::max<size_t>(1, std::min<size_t>(4096, 65536 / n_streams));
    int rc = rnnoise_batch_create(n_streams, d->Qcap, 1, device, &d->rn);
    int prev = 0; cudaGetDevice(&prev); cudaSetDevice(device);
    if (rc == FVAD_OK && (cudaMalloc(&d->d_in, sizeof(float) * d->N * d->Qcap * 480) != cudaSuccess ||
                          cudaMalloc(&d->d_out, sizeof(float) * d->N * d->Qcap * 480) != cudaSuccess ||
                          cudaMalloc(&d->d_vad, sizeof(float) * d->N * d->Qcap) != cudaSuccess))
        rc = set_error(FVAD_EOUT_OF_MEMORY, "device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    cudaSetDevice(prev);
    if (rc != FVAD_OK) { fvad_denoiser_batch_destroy(d); return rc; }
    *out = d;
    return FVAD_OK;
}

extern "C" int fvad_denoiser_batch_run(fvad_denoiser_batch *d, const float *const *pcm, size_t n_frames, float *const *out,
                                       float *const *vad) {
    if (!d || !pcm || !out) return set_error(FVAD_ENOT_INITIALIZED, "denoiser not initialised");
    int prev = 0;
    FVAD_CUDA_TRY(cudaGetDevice(&prev));
    FVAD_CUDA_TRY(cudaSetDevice(d->device));
    const size_t stride = d->Qcap * 480;
    for (size_t f0 = 0; f0 < n_frames; f0 += d->Qcap) {
        const size_t nf = std::min(d->Qcap, n_frames - f0);
        for (size_t s = 0; s < d->N; s++)
            FVAD_CUDA_TRY(cudaMemcpyAsync(d->d_in + s * stride, pcm[s] + f0 * 480, sizeof(float) * nf * 480, cudaMemcpyHostToDevice, 0));
        FVAD_TRY(rnnoise_batch_process(d->rn, d->d_in, stride, 1, nf, d->d_out, stride, 0, d->d_vad, nullptr, d->Qcap, 0, 0));
        for (size_t s = 0; s < d->N; s++) {
            FVAD_CUDA_TRY(cudaMemcpyAsync(out[s] + f0 * 480, d->d_out + s * stride, sizeof(float) * nf * 480, cudaMemcpyDeviceToHost, 0));
            if (vad) FVAD_CUDA_TRY(cudaMemcpyAsync(vad[s] + f0, d->d_vad + s * d->Qcap, sizeof(float) * nf, cudaMemcpyDeviceToHost, 0));
        }
        FVAD_CUDA_TRY(cudaStreamSynchronize(0));
    }
    FVAD_CUDA_TRY(cudaSetDevice(prev));
    return FVAD_OK;
}

extern "C" int fvad_denoiser_create(fvad_denoiser **out) {              // Denoiser.init, Denoiser.zig:22-31
    if (!out) return set_error(FVAD_EINVALID_ARGUMENT, "null out");
    *out = nullptr;
    fvad_denoiser_batch *b = nullptr;
    FVAD_TRY(fvad_denoiser_batch_create(1, -1, &b));
    fvad_denoiser *d = new (std::nothrow) fvad_denoiser();
    if (!d) { fvad_denoiser_batch_destroy(b); return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed"); }
    d->b = b;
    *out = d;
    return FVAD_OK;
}
extern "C" void fvad_denoiser_destroy(fvad_denoiser *d) {               // Denoiser.deinit, Denoiser.zig:33-39
    if (!d) return;
    fvad_denoiser_batch_destroy(d->b);
    delete d;
}
extern "C" int fvad_denoiser_denoise(fvad_denoiser *d, const float *first, size_t n_first, const float *second, size_t n_second,
                                     float *result, size_t n_result, float *vad) {   // Denoiser.denoise, Denoiser.zig:45-66
    if (!d || !d->b) return set_error(FVAD_ENOT_INITIALIZED, "denoiser not initialised");
    if (n_first + n_second != 480) return set_error(FVAD_EINVALID_FRAME_SIZE, "frame of %zu samples, expected 480", n_first + n_second);
    if (n_result != 480) return set_error(FVAD_EINVALID_RESULT_SIZE, "result of %zu samples, expected 480", n_result);
    if (n_first) memcpy(d->in, first, sizeof(float) * n_first);
    if (n_second) memcpy(d->in + n_first, second, sizeof(float) * n_second);
    const float *pin = d->in;
    float v = 0;
    float *pv = &v;
    FVAD_TRY(fvad_denoiser_batch_run(d->b, &pin, 1, &result, &pv));
    if (vad) *vad = v;
    return FVAD_OK;
}
