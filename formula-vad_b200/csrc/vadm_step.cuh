// vadm_step.cuh -- one step of VADMachine.run (src/AudioPipeline/VADMachine.zig:126-310) over RollingAverage
// (src/structures/RollingAverage.zig:34-56), shared by the pipeline's per-tile scan (k_vadm, vadm.cu: rings contiguous per
// machine, state carried between launches) and the whole-run config sweeps (k_vadm_sweep, vadm_sweep.cu: ring entry i of
// scan s at rings[i * n_scans + s], state in registers).  `Ring` is how entry i of a ring is reached.
#pragma once
#include "vadm.cuh"

namespace fvad {

struct RingFlat {                  // contiguous ring
    double *p;
    __device__ __forceinline__ double &operator[](int i) const { return p[i]; }
};
struct RingCol {                   // column of a [ring index][scan] matrix
    double *p;
    size_t pitch;
    __device__ __forceinline__ double &operator[](int i) const { return p[(size_t)i * pitch]; }
};

template <class Ring> __device__ __forceinline__ double ring_avg_exact(const Ring &r, int cnt) {      // RollingAverage.zig:45-56
    double a = 0.0;
    const double scalar = 1.0 / (double)cnt;
    for (int i = 0; i < cnt; i++) a += r[i] * scalar;
    return a;
}

template <class Ring> __device__ __forceinline__ double ring_push_exact(const Ring &r, int len, int &w, int &cnt, float sample) {
    r[w] = (double)sample;                                                                             // RollingAverage.zig:34-43
    if (++w == len) w = 0;
    if (cnt < len) cnt++;
    return ring_avg_exact(r, cnt);
}

template <class Ring> __device__ __forceinline__ double ring_sum_blocked(const Ring &r, int cnt) {    // raw sum, four partial sums
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    int i = 0;
    for (; i + 3 < cnt; i += 4) { a0 += r[i]; a1 += r[i + 1]; a2 += r[i + 2]; a3 += r[i + 3]; }
    for (; i < cnt; i++) a0 += r[i];
    return (a0 + a1) + (a2 + a3);
}

// The long-term ring (4218 entries by default) is re-summed in index order on every push in the reference; its value
// feeds one comparison, `short_term > last_avg * factor` (VADMachine.zig:158-160).  We carry a running RAW sum `rsum` of
// the ring together with a running bound `err` on |rsum - (real sum of the ring)|:
//   full ring   rsum' = fl(rsum + fl(nv - old))   err += 2^-52 (|rsum'| + |nv| + |old|)
//   filling     rsum' = fl(rsum + nv)             err += 2^-52 |rsum'|
//   refresh     rsum  = four-way blocked sum      err  = rsum * (n/4 + 4) * 2^-52
// (every term is twice the rounding error it covers: the bound is ABSOLUTE, so a ring that drains from 1e-3 to 1e-25 keeps
// the error of its large past -- and is refreshed for it, below).  With all entries >= 0 the reference's own value is
// R = fl(1/n) * realsum * (1 + e), |e| <= n * 2^-53, so |fl(fl(rsum * fl(1/n)) * factor) - fl(R * factor)| is at most
// thr * k_rel + err * fl(1/n) * factor * (1 + 1e-6) with k_rel = 1e-9 + n * 2^-52.  Outside that margin the comparison is
// decided by rsum; inside it -- or once a negative / NaN value has entered the ring -- the reference's in-order sum is
// evaluated.  The running sum is refreshed whenever err exceeds 2^-30 * rsum (a large entry leaving the ring cancels the
// sum while the bound keeps its size; plain accumulation needs ~10^6 pushes to get there).  The decision is always the
// reference's.  The entry the next push will replace is fetched right after the previous push (old_next): the load's
// latency (the ring lives in L2) then overlaps the following record instead of sitting between two decisions.
template <class Ring> __device__ __forceinline__ void long_refresh(const Ring &r, LongRing &L) {
    L.rsum = ring_sum_blocked(r, L.cnt);
    L.err = L.rsum * ((double)(L.cnt / 4 + 4) * 0x1p-52);
}

template <class Ring> __device__ __forceinline__ void long_push(const Ring &r, int len, LongRing &L, float sample) {
    const double nv = (double)sample;
    if (!(nv >= 0.0)) L.poisoned = 1;
    if (L.cnt < len) {                                        // still filling (only without initial_long_term_avg)
        r[L.w] = nv;
        L.cnt++;
        L.scalar = 1.0 / (double)L.cnt;
        L.rsum += nv;
        L.err += 0x1p-52 * fabs(L.rsum);
    } else {
        const double old = L.old_next;                        // == r[L.w]
        r[L.w] = nv;
        L.rsum += (nv - old);
        L.err += 0x1p-52 * (fabs(L.rsum) + fabs(nv) + fabs(old));
    }
    if (++L.w == len) L.w = 0;
    if (L.cnt == len) L.old_next = r[L.w];                    // for the next push; nothing writes the ring in between
    L.has_last = 1; L.exact = 0;
    if (!(L.err <= L.rsum * 0x1p-30)) long_refresh(r, L);
}

// `short_term > last_avg * factor` (VADMachine.zig:158-160) with last_avg taken from the running sum where that is beyond
// doubt, from the reference's in-order sum otherwise.  fallbacks counts the latter.
template <class Ring>
__device__ __forceinline__ bool long_exceeds(const Ring &r, LongRing &L, double short_term, double factor, double k_rel,
                                             unsigned long long &fallbacks) {
    if (!L.exact && !L.poisoned) {
        const double thr = (L.rsum * L.scalar) * factor;
        if (L.err == 0.0) return short_term > thr;            // every entry is zero: rsum IS the sum
        const double tol = thr * k_rel + (L.err * L.scalar) * factor * 1.000001;
        if (short_term > thr + tol) return true;
        if (short_term < thr - tol) return false;
    }
    if (!L.exact) { L.last = ring_avg_exact(r, L.cnt); L.exact = 1; fallbacks++; }   // the reference's own sum
    return short_term > L.last * factor;
}

typedef VadmState VadmRegs;        // one machine's state while it scans

// State after VADMachine.init (VADMachine.zig:65-117): rings cleared, the long ring pre-filled with
// initial_long_term_avg when there is one (RollingAverage.zig:19-25).
template <class Ring>
__device__ __forceinline__ void vadm_regs_init(const VadmParams &P, const Ring &rl, const Ring &rs, const Ring &rr, VadmRegs &S) {
    for (int i = 0; i < P.long_len; i++) rl[i] = P.has_initial ? P.initial : 0.0;
    for (int i = 0; i < P.short_len; i++) rs[i] = 0.0;
    for (int i = 0; i < P.ratio_len; i++) rr[i] = 0.0;
    S = VadmRegs{};
    if (P.has_initial) {
        S.L.cnt = P.long_len;
        S.L.scalar = 1.0 / (double)S.L.cnt;
        S.L.last = ring_avg_exact(rl, S.L.cnt);
        S.L.has_last = 1; S.L.exact = 1;
        S.L.poisoned = !(P.initial >= 0.0);
        S.L.old_next = rl[0];
        long_refresh(rl, S.L);
    }
}

struct VadmStepOut {
    int rec;                       // fvad_recording_state of this step (main machine: drives beginCapture / endCapture)
    unsigned long long sample_number;
    bool emit;                     // a segment was completed: `seg`
    fvad_segment seg;
};

// VADMachine.run for one STFT frame: min_volume = min over channels of the speech-band sum, volume_ratio and vad from
// fft_input, index = fft_result.index.
template <class Ring>
__device__ __forceinline__ void vadm_step(const VadmParams &P, double k_rel, const Ring &rl, const Ring &rs, const Ring &rr,
                                          VadmRegs &S, float min_volume, float volume_ratio, float vad,
                                          unsigned long long index, VadmStepOut &out) {
    const double short_term = ring_push_exact(rs, P.short_len, S.short_w, S.short_cnt, min_volume);       // VADMachine.zig:155
    const double cvr = ring_push_exact(rr, P.ratio_len, S.ratio_w, S.ratio_cnt, volume_ratio);            // :156
    bool primary;
    if (!S.L.has_last) {
        const double base = P.has_initial ? P.initial : short_term;                                        // :158
        primary = short_term > base * P.factor;
    } else primary = long_exceeds(rl, S.L, short_term, P.factor, k_rel, S.fallbacks);                      // :159-160
    const bool met = primary && (cvr > P.ratio_threshold);
    if (!met) long_push(rl, P.long_len, S.L, min_volume);                                                  // :165-167

    out.rec = FVAD_REC_NONE; out.sample_number = 0; out.emit = false;
    const int from_state = S.state;
    switch (S.state) {                                                                                     // :176-221
    case 0:   // closed
        if (met) { S.state = 1; S.start_idx = index; }
        break;
    case 1:   // opening
        if (met && index - S.start_idx >= P.min_consecutive) {
            S.state = 2; out.rec = FVAD_REC_STARTED;
            out.sample_number = P.start_buffer > S.start_idx ? 0 : S.start_idx - P.start_buffer;          // :297-302
        } else if (!met) S.state = 0;
        break;
    case 2:   // open
        if (!met) { S.state = 3; S.end_idx = index; }
        break;
    case 3:   // closing
        if (met) S.state = 2;
        else if (index - S.end_idx >= P.max_gap) {
            S.state = 0;
            // onSpeechEnd (VADMachine.zig:252-294)
            const float length_realtime = (float)(S.end_idx - S.start_idx) / P.sample_rate_f;
            if (length_realtime >= P.min_dur) {
                out.seg.sample_from = P.start_buffer > S.start_idx ? 0 : S.start_idx - P.start_buffer;
                out.seg.sample_to = S.end_idx + P.end_buffer;                                             // :305-310
                out.seg.debug_rnn_vad = S.sp_vad / (float)S.sp_vad_cnt;
                out.seg.debug_avg_speech_vol_ratio = S.sp_ratio / (float)S.sp_ratio_cnt;
                out.emit = true;
                out.rec = FVAD_REC_COMPLETED; out.sample_number = S.end_idx + P.end_buffer;
            } else out.rec = FVAD_REC_ABORTED;
        }
        break;
    }
    if (from_state == 0 && S.state == 1) {                                                                 // trackSpeechStats, :232-250
        S.sp_vad = vad; S.sp_vad_cnt = 1; S.sp_ratio = volume_ratio; S.sp_ratio_cnt = 1;
    } else if (from_state == 1 || from_state == 2) {
        S.sp_vad += vad; S.sp_vad_cnt += 1; S.sp_ratio += volume_ratio; S.sp_ratio_cnt += 1;
    }
}

__host__ __device__ inline double vadm_k_rel(int long_len) { return 1e-9 + (double)long_len * 0x1p-52; }

}  // namespace fvad
