// rn_emu.cu -- HOST emulation of the RNNoise kernels (TEST AID, not part of libfvad_cuda.so and
// not a fallback: nothing in the product calls it).  It runs the very same rn_core.cuh phase code
// the CUDA kernels run, lane by lane on the CPU, through the same tile / carry structure as
// rnnoise.cu, so that the warp-level re-organisation can be checked against the oracle without a GPU.
#include "../rn_core.cuh"
#include "../rn_tables.cuh"
#include <cstring>
#include <vector>

using namespace fvad;
using namespace fvad::rn;

namespace {
struct Emu {
    HostTables ht;
    Tables t;
    int C = 1;
    // persistent
    float biq[2] = {0, 0};
    int last_period = 0; float last_gain = 0;
    long long nz_total = 0;
    float lastg[kNB] = {0};
    std::vector<float> rnn_scr;                 // vectors incl. GRU states (persist)
    std::vector<float> hp_hist, ceps_hist, tail_carry;
};
}  // namespace

extern "C" {

void *rn_emu_create(int n_channels) {
    Emu *e = new Emu();
    if (host_tables_build(&e->ht, nullptr) != FVAD_OK) { delete e; return nullptr; }
    e->t = e->ht.view();
    e->C = n_channels;
    e->rnn_scr.assign(kRv_end, 0.f);
    e->hp_hist.assign(kPHist, 0.f);
    e->ceps_hist.assign(8 * kNB, 0.f);
    e->tail_carry.assign(kFrame, 0.f);
    return e;
}
void rn_emu_destroy(void *h) { delete (Emu *)h; }

// One tile: Tn time frames of C channels.  in/out rows: [c][Tn*480] normalised PCM.
// Optional traces per virtual frame q = t*C + c: vad[Q], pitch[Q], silence[Q], feat[Q*42], gains[Q*22], hp[Q*480]
int rn_emu_tile(void *h, const float *in, size_t in_stride, int Tn, float *out, size_t out_stride, float *vad,
                int *pitch, int *silence, float *feat_out, float *gains_out, float *hp_out) {
    Emu &e = *(Emu *)h;
    const Tables &t = e.t;
    const int C = e.C, Q = Tn * C;
    std::vector<float> hp(kPHist + (size_t)Q * kFrame);
    memcpy(hp.data(), e.hp_hist.data(), sizeof(float) * kPHist);
    // k_rn_biquad: sequential per state over the virtual frame sequence
    for (int q = 0; q < Q; q++) {
        const int tt = q / C, c = q % C;
        float sm0[33], sm1[33];
        int flag = 0;
        for (int b0 = 0; b0 < kFrame; b0 += 32)
            biquad_block_warp(&in[c * in_stride + (size_t)tt * kFrame + b0], &hp[kPHist + (size_t)q * kFrame + b0], e.biq[0], e.biq[1], sm0, sm1, &flag);
    }
    std::vector<cpx> X((size_t)Q * kFreq), P((size_t)Q * kFreq);
    std::vector<float> Ex(Q * kNB), Ep(Q * kNB), Exp(Q * kNB), ceps(Q * kNB), featB(Q * 8), feat((size_t)Q * kNFeat, 0.f);
    std::vector<int> sil(Q), pidx(Q);
    std::vector<PitchCand> cand(Q);
    std::vector<float> scr(4096);
    // k_rn_analysis / k_rn_pitch: warp per frame
    for (int q = 0; q < Q; q++) analysis_warp(t, hp.data(), q, &X[(size_t)q * kFreq], &Ex[q * kNB], &ceps[q * kNB], &sil[q], scr.data());
    for (int q = 0; q < Q; q++) {
        const float *xs[1] = {hp.data() + (size_t)q * kFrame};
        PitchCand *outs[1] = {&cand[q]};
        pitch_blk<1>(xs, outs, scr.data());
    }
    // k_rn_pitch_seq
    for (int q = 0; q < Q; q++) {
        float g;
        pidx[q] = pitch_decide(cand[q], e.last_period, e.last_gain, &g);
        e.last_period = pidx[q]; e.last_gain = g;
    }
    // k_rn_pfeat
    for (int q = 0; q < Q; q++)
        pfeat_warp(t, hp.data(), q, pidx[q], &X[(size_t)q * kFreq], &Ex[q * kNB], &P[(size_t)q * kFreq], &Ep[q * kNB], &Exp[q * kNB],
                   &featB[q * 8], scr.data());
    // k_rn_compact: ranks and the compacted cepstrum list with 8 entries of history
    std::vector<int> nzlist; std::vector<long long> rank(Q, -1);
    std::vector<float> cepsc((size_t)(8 + Q) * kNB);
    memcpy(cepsc.data(), e.ceps_hist.data(), sizeof(float) * 8 * kNB);
    for (int q = 0; q < Q; q++)
        if (!sil[q]) {
            rank[q] = e.nz_total + (long long)nzlist.size();
            memcpy(&cepsc[(size_t)(8 + nzlist.size()) * kNB], &ceps[q * kNB], sizeof(float) * kNB);
            nzlist.push_back(q);
        }
    const int nz = (int)nzlist.size();
    // k_rn_cepsfeat
    for (int j = 0; j < nz; j++) {
        const int q = nzlist[j];
        cepsfeat_warp(&cepsc[(size_t)(8 + j) * kNB], rank[q], &featB[q * 8], &feat[(size_t)q * kNFeat], scr.data());
    }
    // k_rn_rnn
    std::vector<float> gains((size_t)Q * kNB, 0.f), vadv(Q, 0.f);
    for (int j = 0; j < nz; j++) {
        const int q = nzlist[j];
        memcpy(e.rnn_scr.data() + kRv_feat, &feat[(size_t)q * kNFeat], sizeof(float) * kNFeat);
        RnnOut<1> out; out.gains[0] = &gains[(size_t)q * kNB]; out.vad[0] = &vadv[q];
        rnn_frame_cta<1>(e.ht.weights2.data(), e.ht.bias.data(), t.tansig, e.rnn_scr.data(), 1u, out);
    }
    // k_rn_gain_scan (denoise.c:478-482)
    std::vector<float> gsm((size_t)Q * kNB, 0.f);
    for (int b = 0; b < kNB; b++)
        for (int q = 0; q < Q; q++)
            if (!sil[q]) {
                const float a = .6f * e.lastg[b];
                float g = gains[(size_t)q * kNB + b];
                g = g > a ? g : a;
                e.lastg[b] = g;
                gsm[(size_t)q * kNB + b] = g;
            }
    // k_rn_synth
    std::vector<float> head((size_t)Q * kFrame), tail((size_t)(Q + 1) * kFrame);
    memcpy(tail.data(), e.tail_carry.data(), sizeof(float) * kFrame);
    for (int q = 0; q < Q; q++)
        synth_warp(t, &X[(size_t)q * kFreq], &P[(size_t)q * kFreq], &Ex[q * kNB], &Ep[q * kNB], &Exp[q * kNB], &gains[(size_t)q * kNB],
                   &gsm[(size_t)q * kNB], sil[q], &head[(size_t)q * kFrame], &tail[(size_t)(q + 1) * kFrame], scr.data());
    // k_rn_out: overlap-add, back to normalised PCM (Denoiser.zig:88-94)
    const float inv = 1.0f / 32767.0f;
    for (int q = 0; q < Q; q++) {
        const int tt = q / C, c = q % C;
        for (int i = 0; i < kFrame; i++) {
            float v = head[(size_t)q * kFrame + i] + tail[(size_t)q * kFrame + i];
            v *= inv;
            out[c * out_stride + (size_t)tt * kFrame + i] = v;
        }
    }
    // carries
    if (Q > 0) {
        memcpy(e.hp_hist.data(), hp.data() + (size_t)Q * kFrame, sizeof(float) * kPHist);
        memcpy(e.ceps_hist.data(), cepsc.data() + (size_t)nz * kNB, sizeof(float) * 8 * kNB);
        memcpy(e.tail_carry.data(), tail.data() + (size_t)Q * kFrame, sizeof(float) * kFrame);
        e.nz_total += nz;
    }
    for (int q = 0; q < Q; q++) {
        if (vad) vad[q] = vadv[q];
        if (pitch) pitch[q] = pidx[q];
        if (silence) silence[q] = sil[q];
        if (feat_out) memcpy(feat_out + (size_t)q * kNFeat, &feat[(size_t)q * kNFeat], sizeof(float) * kNFeat);
        if (gains_out) memcpy(gains_out + (size_t)q * kNB, &gains[(size_t)q * kNB], sizeof(float) * kNB);
        if (hp_out) memcpy(hp_out + (size_t)q * kFrame, &hp[kPHist + (size_t)q * kFrame], sizeof(float) * kFrame);
    }
    return 0;
}

// The protocol of the biquad's chain / verify / fix-up kernels (rnnoise.cu, forms 4 and 5) on the host, one state, n samples
// (a multiple of 32): the lane's fast step in iterations of `it` samples with the denormal test taken four samples before the
// end of the previous iteration, the record (m0 before every sample, m1 before every eighth), then per 32-sample block the
// hazard analysis, the replay of hazardous blocks from the recorded state and the comparison of its end state.  `plant` >= 0
// flips a bit of m0 after that iteration (what knob bq_dbg=32 does).  y = the result after verification and fix-up.
// counts[0] = samples flagged, [1] = blocks replayed, [2] = first block whose replay disagreed (-1: none), [3] = iterations
// that took the double path.
int rn_emu_biquad_protocol(const float *in, size_t n, int it, long plant, float *y, float *state, long *counts) {
    if (n % 32 || it % 8 || it <= 0) return -1;
    std::vector<float> M0(n + 1), M1(n / 8 + 1), xs(n);
    for (size_t i = 0; i < n; i++) xs[i] = in[i] * 32767.0f;
    float m0 = state[0], m1 = state[1];
    const float s0 = m0, s1 = m1;
    auto is_tiny = [](float a, float b) {
        a = a < 0 ? -a : a; b = b < 0 ? -b : b;
        return (a != 0.0f && a < 1.0e-25f) || (b != 0.0f && b < 1.0e-25f);
    };
    bool tiny = is_tiny(m0, m1);
    long n_tiny_it = 0, h = 0;
    for (size_t i0 = 0; i0 < n; i0 += it, h++) {
        const size_t len = i0 + it <= n ? it : n - i0;
        bool next_tiny = tiny;
        if (tiny) n_tiny_it++;
        for (size_t i = 0; i < len; i++) {
            if ((i0 + i) % 8 == 0) M1[(i0 + i) / 8] = m1;
            M0[i0 + i] = m0;
            if (tiny) biquad_step_x(xs[i0 + i], m0, m1);
            else biquad_chain_step(xs[i0 + i], m0, m1);
            if (!tiny && i + 1 == len - 4) next_tiny = is_tiny(m0, m1);
        }
        tiny = tiny ? is_tiny(m0, m1) : next_tiny;
        if (h == plant) m0 = u2f(f2u(m0) ^ 0x1000u);
    }
    M0[n] = m0; M1[n / 8] = m1;
    (void)s0; (void)s1;
    long flagged = 0, replayed = 0, broken = -1;
    for (size_t b0 = 0; b0 < n; b0 += 32) {
        bool any = false;
        for (int l = 0; l < 32; l++) {
            const size_t t = b0 + l;
            float a = M0[t], b;
            if (t % 8 == 0) b = M1[t / 8];
            else b = fma_rn(-0.99600f, xs[t - 1] + M0[t - 1], xs[t - 1]);
            bool hz = false;
            const float yv = biquad_step_eft_x(xs[t], a, b, hz);
            hz |= f2u(a) != f2u(M0[t + 1]);
            const float bl = (t + 1) % 8 == 0 ? M1[(t + 1) / 8] : fma_rn(-0.99600f, yv, xs[t]);
            hz |= f2u(b) != f2u(bl);
            y[t] = yv;
            if (hz) { flagged++; any = true; }
        }
        if (any) {
            replayed++;
            float r0 = M0[b0], r1 = M1[b0 / 8];
            for (int l = 0; l < 32; l++) y[b0 + l] = biquad_step_x(xs[b0 + l], r0, r1);
            if ((f2u(r0) != f2u(M0[b0 + 32]) || f2u(r1) != f2u(M1[(b0 + 32) / 8])) && broken < 0) broken = (long)(b0 / 32);
        }
    }
    if (broken >= 0) {                                         // k_rn_biquad_fixup
        size_t t0 = (size_t)broken * 32;
        m0 = M0[t0]; m1 = M1[t0 / 8];
        for (size_t t = t0; t < n; t++) y[t] = biquad_step_x(xs[t], m0, m1);
    }
    state[0] = m0; state[1] = m1;
    counts[0] = flagged; counts[1] = replayed; counts[2] = broken; counts[3] = n_tiny_it;
    return 0;
}
}
