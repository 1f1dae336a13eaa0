// vadm.cu -- see vadm.cuh.
//
// The reference recomputes every rolling mean by re-summing its whole ring in index
// order in f64 on every push (RollingAverage.zig:45-56): 4218 dependent adds per STFT
// frame for the long-term ring.  Its value feeds exactly one thing, the comparison
// `short_term > last_avg * factor` (VADMachine.zig:158-160).  We keep a running value of
// the same sum whose error is bounded far below 1e-9 relative, decide with it when the
// comparison is not within 1e-9 of a tie, and otherwise redo the reference's exact
// in-order sum.  The decision -- hence the segment list -- is therefore always the
// reference's.  The two short rings (4 and 11 entries by default) are summed exactly.
#include "vadm.cuh"
#include <cmath>
#include <cstring>

namespace fvad {

int vadm_params_from_config(const fvad_vad_machine_config &c, size_t sample_rate, size_t fft_size, VadmParams *p,
                            size_t *ring_doubles) {
    memset(p, 0, sizeof *p);
    const float sr = (float)sample_rate;
    const float eval_per_sec = sr / (float)fft_size;                                  // VADMachine.zig:70
    size_t long_len = (size_t)(eval_per_sec * c.long_term_speech_avg_sec);            // :71
    size_t short_len = (size_t)(eval_per_sec * c.short_term_speech_avg_sec);          // :72
    size_t ratio_len = (size_t)(eval_per_sec * c.channel_vol_ratio_avg_sec);          // :73
    if (ratio_len == 0)
        return set_error(FVAD_EINVALID_ARGUMENT, "channel_vol_ratio_avg_sec gives an empty ring (the reference would index out of bounds)");
    if (long_len < 1) long_len = 1;                                                   // @max(1, ..) :77,84
    if (short_len < 1) short_len = 1;
    if (long_len > (1u << 24) || short_len > (1u << 16) || ratio_len > (1u << 16))
        return set_error(FVAD_EINVALID_ARGUMENT, "rolling-average window too long");
    p->long_len = (int)long_len; p->short_len = (int)short_len; p->ratio_len = (int)ratio_len;
    p->has_initial = c.has_initial_long_term_avg != 0;
    p->initial = c.initial_long_term_avg;
    p->factor = (double)c.speech_threshold_factor;
    p->ratio_threshold = (double)c.channel_vol_ratio_threshold;
    p->min_consecutive = (unsigned long long)(sr * c.min_consecutive_sec_to_open);    // :150
    p->max_gap = (unsigned long long)(sr * c.max_speech_gap_sec);                     // :152
    p->start_buffer = (unsigned long long)(sr * 2);                                   // :299
    p->end_buffer = (unsigned long long)(sr * 2);                                     // :307
    p->min_dur = c.min_vad_duration_sec;
    p->sample_rate_f = sr;
    *ring_doubles = long_len + short_len + ratio_len;
    return FVAD_OK;
}

void vadm_init_host(const VadmParams &p, VadmState *st, double *rings) {
    memset(st, 0, sizeof *st);
    const size_t n = (size_t)p.long_len + p.short_len + p.ratio_len;
    for (size_t i = 0; i < n; i++) rings[i] = 0.0;
    if (p.has_initial) {                                                              // RollingAverage.zig:19-25
        for (int i = 0; i < p.long_len; i++) rings[i] = p.initial;
        st->long_cnt = p.long_len;
        double a = 0.0, scalar = 1.0 / (double)p.long_len;
        for (int i = 0; i < p.long_len; i++) a += rings[i] * scalar;
        st->long_last = a; st->approx = a; st->long_has_last = 1; st->long_exact = 1;
    }
}

__device__ __forceinline__ double ring_avg_exact(const double *ring, int cnt) {      // RollingAverage.zig:45-56
    double a = 0.0;
    const double scalar = 1.0 / (double)cnt;
    for (int i = 0; i < cnt; i++) a += ring[i] * scalar;
    return a;
}

__device__ __forceinline__ double ring_sum_blocked(const double *ring, int cnt) {    // refresh of the running value
    const double scalar = 1.0 / (double)cnt;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    int i = 0;
    for (; i + 3 < cnt; i += 4) {
        a0 += ring[i] * scalar; a1 += ring[i + 1] * scalar; a2 += ring[i + 2] * scalar; a3 += ring[i + 3] * scalar;
    }
    for (; i < cnt; i++) a0 += ring[i] * scalar;
    return (a0 + a1) + (a2 + a3);
}

__device__ __forceinline__ double ring_push_exact(double *ring, int len, int &w, int &cnt, float sample) {
    ring[w] = (double)sample;                                                         // RollingAverage.zig:34-43
    w = (w + 1) % len;
    if (cnt < len) cnt++;
    return ring_avg_exact(ring, cnt);
}

__global__ void k_vadm(const VadmParams *__restrict__ params, VadmState *__restrict__ states, double *__restrict__ rings,
                       size_t n_streams, int n_machines, VadmIo io) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_streams * (size_t)n_machines) return;
    const size_t s = id / n_machines;
    const int k = (int)(id - s * n_machines);
    const VadmParams P = params[k];
    VadmState S = states[id];
    // ring pool: per stream a block holding every machine's rings
    size_t per_stream = 0;
    {
        const VadmParams last = params[n_machines - 1];
        per_stream = last.ring_off + (size_t)last.long_len + last.short_len + last.ratio_len;
    }
    double *ring_long = rings + s * per_stream + P.ring_off;
    double *ring_short = ring_long + P.long_len;
    double *ring_ratio = ring_short + P.short_len;
    const double kTie = 1e-9;
    const int C = io.n_channels;

    for (size_t f = 0; f < io.n_frames; f++) {
        // min over channels of the speech-band sum (VADMachine.zig:134-146)
        float min_volume = 999.0f;
        for (int c = 0; c < C; c++) {
            float v = io.band[((size_t)P.band * io.band_rows + s * C + c) * io.band_frame_stride + f];
            if (v < min_volume) min_volume = v;
        }
        const float volume_ratio = io.vol_ratio[s * io.ratio_stride + f];
        const float vad = io.vad ? io.vad[s * io.ratio_stride + f] : 0.0f;
        const unsigned long long index = io.first_index + f * io.index_step;

        const double short_term = ring_push_exact(ring_short, P.short_len, S.short_w, S.short_cnt, min_volume);
        const double cvr = ring_push_exact(ring_ratio, P.ratio_len, S.ratio_w, S.ratio_cnt, volume_ratio);

        bool primary;
        if (!S.long_has_last) {
            const double base = P.has_initial ? P.initial : short_term;               // VADMachine.zig:158
            primary = short_term > base * P.factor;
        } else {
            bool decided = false;
            if (!S.long_exact) {
                const double thr = S.approx * P.factor;
                if (thr > 0.0 && thr < 1e300) {
                    if (short_term > thr * (1.0 + kTie)) { primary = true; decided = true; }
                    else if (short_term < thr * (1.0 - kTie)) { primary = false; decided = true; }
                }
                if (!decided) {                                                       // near a tie: the reference's own sum
                    S.long_last = ring_avg_exact(ring_long, S.long_cnt);
                    S.long_exact = 1;
                    S.exact_fallbacks++;
                }
            }
            if (!decided) primary = short_term > S.long_last * P.factor;
        }
        const bool met = primary && (cvr > P.ratio_threshold);

        if (!met) {                                                                   // VADMachine.zig:165-167
            const double nv = (double)min_volume;
            if (S.long_cnt < P.long_len) {
                // ring still filling (only without initial_long_term_avg): scalar changes each push
                S.long_last = ring_push_exact(ring_long, P.long_len, S.long_w, S.long_cnt, min_volume);
                S.approx = S.long_last; S.long_exact = 1; S.long_has_last = 1; S.since_refresh = 0;
            } else {
                const double scalar = 1.0 / (double)P.long_len;
                const double old = ring_long[S.long_w];
                ring_long[S.long_w] = nv;
                S.long_w = (S.long_w + 1) % P.long_len;
                S.approx += (nv * scalar - old * scalar);
                S.long_exact = 0; S.long_has_last = 1;
                if (++S.since_refresh >= 512 || !(S.approx >= 0.0)) {
                    S.approx = ring_sum_blocked(ring_long, P.long_len);
                    S.since_refresh = 0;
                }
            }
        }

        int rec = FVAD_REC_NONE;
        unsigned long long sample_number = 0;
        const int from_state = S.state;
        bool emit = false;
        fvad_segment seg;
        switch (S.state) {                                                            // VADMachine.zig:176-221
        case 0:   // closed
            if (met) { S.state = 1; S.start_idx = index; }
            break;
        case 1: { // opening
            const unsigned long long since = index - S.start_idx;
            if (met && since >= P.min_consecutive) {
                S.state = 2; rec = FVAD_REC_STARTED;
                sample_number = P.start_buffer > S.start_idx ? 0 : S.start_idx - P.start_buffer;
            } else if (!met) S.state = 0;
            break; }
        case 2:   // open
            if (!met) { S.state = 3; S.end_idx = index; }
            break;
        case 3: { // closing
            const unsigned long long since = index - S.end_idx;
            if (met) S.state = 2;
            else if (since >= P.max_gap) {
                S.state = 0;
                // onSpeechEnd (VADMachine.zig:252-294)
                const unsigned long long from = S.start_idx, to = S.end_idx;
                const float length_realtime = (float)(to - from) / P.sample_rate_f;
                const bool dur_met = length_realtime >= P.min_dur;
                if (dur_met) {
                    seg.sample_from = P.start_buffer > from ? 0 : from - P.start_buffer;
                    seg.sample_to = to + P.end_buffer;
                    seg.debug_rnn_vad = S.sp_vad / (float)S.sp_vad_cnt;
                    seg.debug_avg_speech_vol_ratio = S.sp_ratio / (float)S.sp_ratio_cnt;
                    emit = true;
                    rec = FVAD_REC_COMPLETED; sample_number = to + P.end_buffer;
                } else { rec = FVAD_REC_ABORTED; sample_number = 0; }
            }
            break; }
        }
        // trackSpeechStats (VADMachine.zig:232-250)
        if (from_state == 0 && S.state == 1) {
            S.sp_vad = vad; S.sp_vad_cnt = 1; S.sp_ratio = volume_ratio; S.sp_ratio_cnt = 1;
        } else if (from_state == 1 || from_state == 2) {
            S.sp_vad += vad; S.sp_vad_cnt += 1; S.sp_ratio += volume_ratio; S.sp_ratio_cnt += 1;
        }
        if (emit) {
            unsigned slot = atomicAdd(io.seg_count, 1u);
            if (slot < io.seg_cap) {
                SegRecord r; r.stream = (unsigned)s; r.machine = (unsigned)k; r.seg = seg;
                io.segs[slot] = r;
            }
        }
        if (k == 0) {
            if (rec != FVAD_REC_NONE) {
                unsigned slot = atomicAdd(io.event_count, 1u);
                if (slot < io.event_cap) {
                    EventRecord e; e.stream = (unsigned)s; e._pad = 0;
                    e.ev.at_index = index; e.ev.recording_state = rec; e.ev._pad = 0; e.ev.sample_number = sample_number;
                    io.events[slot] = e;
                }
            }
            if (io.trace) {
                fvad_stft_record tr;
                tr.index = index; tr.min_volume = min_volume; tr.vol_ratio = volume_ratio; tr.rnn_vad = vad;
                tr.has_vad = io.vad ? 1.0f : 0.0f;
                io.trace[s * io.trace_stride + io.trace_offset + f] = tr;
            }
        }
    }
    states[id] = S;
}

int vadm_launch(const VadmParams *d_params, VadmState *d_states, double *d_rings, size_t n_streams, int n_machines,
                const VadmIo &io, cudaStream_t stream) {
    if (io.n_frames == 0) return FVAD_OK;
    const size_t n = n_streams * (size_t)n_machines;
    const unsigned block = 32;                       // one warp per CTA spreads the scans over the SMs
    FVAD_LAUNCH(k_vadm, (unsigned)ceil_div(n, block), block, 0, stream, d_params, d_states, d_rings, n_streams,
                n_machines, io);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad
