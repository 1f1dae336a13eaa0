// vadm.cu -- see vadm.cuh.
//
// The reference recomputes every rolling mean by re-summing its whole ring in index
// order in f64 on every push (RollingAverage.zig:45-56): 4218 dependent adds per STFT
// frame for the long-term ring.  Its value feeds exactly one thing, the comparison
// `short_term > last_avg * factor` (VADMachine.zig:158-160).  We keep a running sum with a
// running ABSOLUTE error bound, decide with it when the comparison is further from a tie than
// that bound allows, and otherwise redo the reference's exact in-order sum (vadm_step.cuh).
// The decision -- hence the segment list -- is therefore always the reference's.  The two
// short rings (4 and 11 entries by default) are summed exactly.
#include "vadm_step.cuh"
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

// One thread per (stream, machine): the state VADMachine.init leaves behind.
__global__ void k_vadm_init(const VadmParams *__restrict__ params, VadmState *__restrict__ states, double *__restrict__ rings,
                            size_t n_streams, int n_machines) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_streams * (size_t)n_machines) return;
    const size_t s = id / n_machines;
    const VadmParams P = params[id - s * n_machines];
    const VadmParams last = params[n_machines - 1];
    const size_t per_stream = last.ring_off + (size_t)last.long_len + last.short_len + last.ratio_len;
    double *base = rings + s * per_stream + P.ring_off;
    const RingFlat rl{base}, rs{base + P.long_len}, rr{base + P.long_len + P.short_len};
    VadmState S;
    vadm_regs_init(P, rl, rs, rr, S);
    states[id] = S;
}

int vadm_init_launch(const VadmParams *d_params, VadmState *d_states, double *d_rings, size_t n_streams, int n_machines,
                     cudaStream_t stream) {
    const size_t n = n_streams * (size_t)n_machines;
    if (!n) return FVAD_OK;
    FVAD_LAUNCH(k_vadm_init, (unsigned)ceil_div(n, 32), 32, 0, stream, d_params, d_states, d_rings, n_streams, n_machines);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
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
    const VadmParams last = params[n_machines - 1];
    const size_t per_stream = last.ring_off + (size_t)last.long_len + last.short_len + last.ratio_len;
    double *base = rings + s * per_stream + P.ring_off;
    const RingFlat rl{base}, rs{base + P.long_len}, rr{base + P.long_len + P.short_len};
    const double k_rel = vadm_k_rel(P.long_len);
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
        VadmStepOut o;
        vadm_step(P, k_rel, rl, rs, rr, S, min_volume, volume_ratio, vad, index, o);
        if (o.emit) {
            unsigned slot = atomicAdd(io.seg_count, 1u);
            if (slot < io.seg_cap) {
                SegRecord r; r.stream = (unsigned)s; r.machine = (unsigned)k; r.seg = o.seg;
                io.segs[slot] = r;
            }
        }
        if (k == 0) {
            if (o.rec != FVAD_REC_NONE) {
                unsigned slot = atomicAdd(io.event_count, 1u);
                if (slot < io.event_cap) {
                    EventRecord e; e.stream = (unsigned)s; e._pad = 0;
                    e.ev.at_index = index; e.ev.recording_state = o.rec; e.ev._pad = 0; e.ev.sample_number = o.sample_number;
                    io.events[slot] = e;
                }
            }
            if (io.trace) {
                fvad_stft_record tr;
                tr.index = index; tr.min_volume = min_volume; tr.vol_ratio = volume_ratio; tr.rnn_vad = vad;
                tr.has_vad = io.vad ? 1.0f : 0.0f;
                io.trace[s * io.trace_stride + io.trace_offset + f] = tr;
            }
            if (io.feat_min) {                                    // the run's features kept on the device (fvad_batch_sweep)
                const size_t at = s * io.feat_stride + io.feat_offset + f;
                io.feat_ratio[at] = volume_ratio;
                io.feat_vad[at] = vad;
                for (int bnd = 0; bnd < io.feat_bands; bnd++) {
                    float mv = 999.0f;
                    for (int c = 0; c < C; c++) {
                        const float v = io.band[((size_t)bnd * io.band_rows + s * C + c) * io.band_frame_stride + f];
                        if (v < mv) mv = v;
                    }
                    io.feat_min[(size_t)bnd * io.feat_band_stride + at] = mv;
                }
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
