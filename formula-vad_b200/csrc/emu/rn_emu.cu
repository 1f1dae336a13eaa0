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
}
