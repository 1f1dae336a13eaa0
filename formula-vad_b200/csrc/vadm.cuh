// vadm.cuh -- VADMachine (src/AudioPipeline/VADMachine.zig) as a per-(stream, machine)
// sequential scan over STFT-frame records.
#pragma once
#include "common.cuh"

namespace fvad {

struct VadmParams {                 // derived once from VADMachine.Config (VADMachine.zig:65-117,126-152)
    int band;                       // which band-sum plane this machine reads
    int long_len, short_len, ratio_len;
    int has_initial;
    double initial;
    double factor;                  // speech_threshold_factor widened to f64 (VADMachine.zig:159)
    double ratio_threshold;         // channel_vol_ratio_threshold widened to f64 (VADMachine.zig:160)
    unsigned long long min_consecutive, max_gap, start_buffer, end_buffer;
    float min_dur, sample_rate_f;
    size_t ring_off;                // offset (doubles) of this machine's three rings in the ring pool
};

struct VadmState {
    int state;                      // SpeechState
    int long_w, long_cnt, short_w, short_cnt, ratio_w, ratio_cnt;
    int long_has_last;              // last_avg != null
    int long_exact;                 // long_last holds the exact in-order sum (else only `approx` is current)
    int since_refresh;
    double long_last;               // exact last_avg when long_exact
    double approx;                  // running value of the same sum, error << 1e-9 relative
    unsigned long long start_idx, end_idx;
    float sp_vad, sp_ratio;
    unsigned sp_vad_cnt, sp_ratio_cnt;
    unsigned long long exact_fallbacks;   // statistics: how often the exact re-sum was needed
};

struct SegRecord {                  // appended with an atomic counter; host sorts by (stream, machine, order)
    unsigned stream, machine;
    fvad_segment seg;
};
struct EventRecord {
    unsigned stream, _pad;
    fvad_capture_event ev;
};

struct VadmIo {
    // inputs
    const float *band;              // [band][row = stream*C + c][frame_stride] speech-band sums
    size_t band_rows, band_frame_stride;
    const float *vol_ratio;         // [stream][ratio_stride]
    const float *vad;               // [stream][ratio_stride] or nullptr (fft_input.vad == null)
    size_t ratio_stride;
    int n_channels;
    unsigned long long first_index; // fft_result.index of frame 0
    unsigned long long index_step;  // fft_size
    size_t n_frames;
    // outputs
    SegRecord *segs; unsigned *seg_count; unsigned seg_cap;
    EventRecord *events; unsigned *event_count; unsigned event_cap;
    fvad_stft_record *trace;        // optional [stream][trace_stride], written by machine 0
    size_t trace_stride, trace_offset;
};

int vadm_params_from_config(const fvad_vad_machine_config &c, size_t sample_rate, size_t fft_size, VadmParams *out,
                            size_t *ring_doubles);
// initial state + ring contents for one machine (host side, uploaded once)
void vadm_init_host(const VadmParams &p, VadmState *st, double *rings);
int vadm_launch(const VadmParams *d_params, VadmState *d_states, double *d_rings, size_t n_streams, int n_machines,
                const VadmIo &io, cudaStream_t stream);

}  // namespace fvad
