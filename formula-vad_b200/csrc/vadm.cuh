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

// The long-term ring's running sum with its error bound (vadm_step.cuh).
struct LongRing {
    int w, cnt;
    int has_last, exact, poisoned;  // last_avg != null; `last` holds the exact in-order value; a value < 0 or NaN went in
    double last, rsum, err, scalar; // scalar = 1 / cnt
    double old_next;                // ring[w] fetched ahead of the push that will replace it (valid while cnt == len)
};

struct VadmState {                  // one machine's state: carried between launches by k_vadm, in registers while it scans
    int state;                      // SpeechState: 0 closed, 1 opening, 2 open, 3 closing
    int short_w, short_cnt, ratio_w, ratio_cnt;
    LongRing L;
    unsigned long long start_idx, end_idx;
    float sp_vad, sp_ratio;
    unsigned sp_vad_cnt, sp_ratio_cnt;
    unsigned long long fallbacks;   // statistics: how often the reference's in-order long sum was needed
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
    // optional: the features kept on the device for the whole run (fvad_batch_sweep), written by machine 0:
    // feat_min[band][stream][feat_stride] = min over channels of band `band`, feat_ratio / feat_vad [stream][feat_stride]
    float *feat_min, *feat_ratio, *feat_vad;
    size_t feat_stride, feat_offset, feat_band_stride;
    int feat_bands;
};

int vadm_params_from_config(const fvad_vad_machine_config &c, size_t sample_rate, size_t fft_size, VadmParams *out,
                            size_t *ring_doubles);
// State after VADMachine.init for every (stream, machine): rings cleared / pre-filled, on the device.
int vadm_init_launch(const VadmParams *d_params, VadmState *d_states, double *d_rings, size_t n_streams, int n_machines,
                     cudaStream_t stream);
int vadm_launch(const VadmParams *d_params, VadmState *d_states, double *d_rings, size_t n_streams, int n_machines,
                const VadmIo &io, cudaStream_t stream);

// ---- whole-run sweeps: one scan per (stream, config) over feature planes, state in registers, rings in a
// [ring index][scan] matrix (vadm_sweep.cu) ---------------------------------------------------------------------------
struct VadmSweepIo {
    const float *min_volume;        // [band][stream][rec_stride]: min over channels of the band sum; a config reads plane params.band
    size_t band_stride;             // floats between two band planes
    const float *vol_ratio;         // [stream][rec_stride]
    const float *vad;               // [stream][rec_stride] or nullptr (fft_input.vad == null)
    size_t rec_stride, n_records;
    unsigned long long first_index, index_step;   // fft_result.index of record f = first_index + f * index_step
    fvad_segment *segs;             // [scan][max_segs], scan = stream * n_configs + config, in emission order
    size_t max_segs;
    uint32_t *counts;               // [scan] segments found (may exceed max_segs: the list is truncated)
    unsigned long long *fallbacks;  // optional [scan]
};
// rows of the ring matrix the configs need: max long + max short + max ratio length
void vadm_sweep_rows(const VadmParams *params, size_t n_configs, int *max_long, int *max_short, int *max_ratio);
// d_rings: (max_long + max_short + max_ratio) x (n_streams * n_configs) doubles, contents irrelevant (rewritten first)
int vadm_sweep_launch(const VadmParams *d_params, double *d_rings, int max_long, int max_short, size_t n_streams, int n_configs,
                      const VadmSweepIo &io, cudaStream_t stream);

}  // namespace fvad
