/*
 * fvad.h -- C ABI of libfvad_cuda.so: the B200 (sm_100a) implementation of
 * Formula-VAD's per-frame AudioPipeline hot path.
 *
 * This is the drop-in boundary.  Every entry point names the reference interface
 * it replaces (file:line under the reference tree).  Signatures use plain
 * pointers and sizes only.  There is NO CPU fallback: every compute entry point
 * returns FVAD_ECUDA when no usable CUDA device / kernel image is present.
 *
 * The reference processes ONE stream per thread, one 480-sample frame at a time
 * (src/simulator.zig:217-228).  The GPU is efficient only when many streams x many
 * frames are submitted at once, so the primary object is a BATCH of independent
 * pipelines that share one configuration; the single-pipeline API is a batch of 1.
 *
 * Threading: a handle is owned by one thread at a time (as a reference pipeline
 * is, src/simulator/SimulationInstance.zig:134-142); distinct handles may be used
 * from distinct threads.  Errors: 0 = ok, negative fvad_status; fvad_last_error()
 * gives a thread-local message.
 */
#ifndef FVAD_H
#define FVAD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes: the reference's Zig error names ------- */
typedef enum fvad_status {
    FVAD_OK = 0,
    FVAD_EINVALID_SAMPLE_RATE   = -1,   /* VAD.zig:101-104 error.InvalidSampleRate */
    FVAD_EINVALID_FFT_SIZE      = -2,   /* FFT.zig:29-31 error.InvalidFFTSize (+ sizes this build does not cover) */
    FVAD_EINVALID_SAMPLES_LEN   = -3,   /* FFT.zig:71-73 error.InvalidSamplesLength */
    FVAD_EINVALID_WINDOW_LEN    = -4,   /* FFT.zig:75-77 error.InvalidWindowLength */
    FVAD_EINVALID_RESULT_LEN    = -5,   /* FFT.zig:79-81 error.InvalidResultLength */
    FVAD_EOUT_OF_RANGE          = -6,   /* FFT.zig:121-123,135-137 error.OutOfRange */
    FVAD_ENEGATIVE_FREQUENCY    = -7,   /* FFT.zig:125-127 error.NegativeFrequency */
    FVAD_EINVALID_FRAME_SIZE    = -8,   /* Denoiser.zig:48-50 error.InvalidFrameSize */
    FVAD_EINVALID_RESULT_SIZE   = -9,   /* Denoiser.zig:52-54 error.InvalidResultSize */
    FVAD_ENOT_INITIALIZED       = -10,  /* Denoiser.zig:46 error.NotInitialized */
    FVAD_ECHANNEL_COUNT_MISMATCH= -11,  /* SegmentWriter.zig:63-65 error.ChannelCountMismatch */
    FVAD_EOUT_OF_MEMORY         = -12,  /* error.OutOfMemory (host or device) */
    FVAD_EINVALID_ARGUMENT      = -13,  /* null handle, bad index, zero-length ring, ... */
    FVAD_EUNSUPPORTED           = -14,  /* valid in the reference, not covered by this build */
    FVAD_ECUDA                  = -15,  /* CUDA runtime / driver failure, or no sm_100 device: there is no CPU path */
    FVAD_EINSUFFICIENT_SAMPLES  = -16   /* Spectrogram.zig:34-36 error.InsufficientSamples */
} fvad_status;

const char *fvad_last_error(void);
const char *fvad_status_name(int status);
const char *fvad_version(void);
/* Number of CUDA devices visible (<0: FVAD_ECUDA). */
int fvad_device_count(void);
/* Kernels launched by this library in this process so far (all handles, all threads). */
uint64_t fvad_launch_count(void);
/* Per-kernel timing for bench.py: when enabled, every launch is bracketed by CUDA events on its own
 * stream; fvad_profile_report writes {"kernel": {"launches": n, "total_ms": t}, ...} (JSON) for
 * everything recorded since the previous report and returns the length needed. */
void fvad_profile_enable(int on);
int  fvad_profile_report(char *buf, size_t cap);

/* Measurement / test knobs -- kernel shapes ("bq_block", "bq_form", "biquad_lanes", "rnn_r", "rnn_g", "pitch_f"), tiling
 * ("tile_frames", "slots"), scheduling ("prio", "latency_mode", "serial_stages"), "timeline", "model_blob".  They
 * take effect for handles created afterwards.  value NULL clears the key; key NULL clears all.  This setter is the ONLY
 * way to change them: the library does not read the process environment.  No counterpart in the reference. */
int  fvad_debug_set(const char *key, const char *value);
/* Test aid: out[i] = (float)log10(1e-2 + ex[i]) as the DEVICE evaluates it inside rnnoise's feature extraction
 * (lib/rnnoise/src/denoise.c:349); host arrays.  Lets a test sweep the one libm call of the path whose device and host
 * implementations are not both correctly rounded. */
int  fvad_debug_log10(const float *ex, float *out, size_t n);

/* ---- configuration PODs ---------------------------------------------------------- */
/* VADMachine.Config, src/AudioPipeline/VADMachine.zig:18-39 (same fields, same defaults
 * via fvad_default_config). */
typedef struct fvad_vad_machine_config {
    float  speech_min_freq;
    float  speech_max_freq;
    float  long_term_speech_avg_sec;
    int32_t has_initial_long_term_avg;      /* the Zig field is ?f64; 0 = null */
    double initial_long_term_avg;
    float  short_term_speech_avg_sec;
    float  speech_threshold_factor;
    float  channel_vol_ratio_avg_sec;
    float  channel_vol_ratio_threshold;
    float  min_consecutive_sec_to_open;
    float  max_speech_gap_sec;
    float  min_vad_duration_sec;
} fvad_vad_machine_config;

/* VAD.Config, src/AudioPipeline/VAD.zig:17-23 */
typedef struct fvad_vad_config {
    size_t  fft_size;                                        /* default 2048 */
    int32_t use_denoiser;                                    /* default 1 */
    fvad_vad_machine_config vad_machine_config;
    const fvad_vad_machine_config *alt_vad_machine_configs;  /* may be NULL */
    size_t  n_alt_vad_machine_configs;
} fvad_vad_config;

/* AudioPipeline.Config, src/AudioPipeline.zig:20-26 */
typedef struct fvad_config {
    size_t  sample_rate;
    size_t  n_channels;
    size_t  buffer_length;       /* ?usize; 0 = null (sample_rate*10).  Host ring only: no effect on results */
    fvad_vad_config vad_config;
    int32_t skip_processing;
} fvad_config;

void fvad_default_vad_machine_config(fvad_vad_machine_config *c);
void fvad_default_config(fvad_config *c, size_t sample_rate, size_t n_channels);

/* VAD.VADSpeechSegment, src/AudioPipeline/VAD.zig:25-30 */
typedef struct fvad_segment {
    uint64_t sample_from;
    uint64_t sample_to;
    float    debug_rnn_vad;
    float    debug_avg_speech_vol_ratio;
} fvad_segment;

/* VAD.VADMachineResult, src/AudioPipeline/VAD.zig:32-42, as emitted by the main machine;
 * these are what drive AudioPipeline.beginCapture / endCapture (VAD.zig:362-373). */
typedef enum fvad_recording_state { FVAD_REC_NONE = 0, FVAD_REC_STARTED = 1, FVAD_REC_COMPLETED = 2, FVAD_REC_ABORTED = 3 } fvad_recording_state;
typedef struct fvad_capture_event {
    uint64_t at_index;        /* fft_result.index of the STFT frame that produced it */
    int32_t  recording_state; /* fvad_recording_state */
    int32_t  _pad;
    uint64_t sample_number;
} fvad_capture_event;

/* What VADMachine.run consumes per STFT frame ("shared features", VAD.zig:355-381). */
typedef struct fvad_stft_record {
    uint64_t index;        /* fft_result.index */
    float    min_volume;   /* min over channels of the speech-band sum for the MAIN machine's band */
    float    vol_ratio;    /* fft_input.volume_ratio */
    float    rnn_vad;      /* fft_input.vad orelse 0 */
    float    has_vad;      /* 1.0 if fft_input.vad != null */
} fvad_stft_record;

/* ---- batch of pipelines (primary API) --------------------------------------------- */
typedef struct fvad_batch fvad_batch;

/* AudioPipeline.init x n_streams (src/AudioPipeline.zig:39-75, VAD.zig:97-192), all on CUDA
 * device `device` (-1 = current).  Errors as the reference: sample_rate != 48000 ->
 * FVAD_EINVALID_SAMPLE_RATE; odd/zero fft_size -> FVAD_EINVALID_FFT_SIZE.  fft_size: 2048 runs the specialised
 * bandwidth kernel; every other even size whose half is made of the factors 2, 3, 4, 5 (256, 512, 960, 1024, 1920, 4096,
 * ...) runs KissFFT's generic plan; a half with a larger prime factor -> FVAD_EUNSUPPORTED (kf_bfly_generic is not
 * implemented); with the denoiser fft_size <= 6240 (the 480 -> fft_size re-framing history). */
int  fvad_batch_create(const fvad_config *cfg, size_t n_streams, int device, fvad_batch **out);
/* AudioPipeline.deinit */
void fvad_batch_destroy(fvad_batch *b);

/* AudioPipeline.pushSamples (src/AudioPipeline.zig:86-120) for every stream at once, HOST
 * buffers: pcm[stream * n_channels + channel] points at n_samples planar f32 in [-1,1].
 * Input is borrowed until the call returns.  *first_index = total samples written before. */
int  fvad_batch_push(fvad_batch *b, const float *const *pcm, size_t n_samples, uint64_t *first_index);
/* Same with DEVICE-resident input: sample i of (stream s, channel c) at
 * d_pcm[s*stream_stride + c*channel_stride + i].  Asynchronous on the batch's CUDA streams;
 * the input must stay valid until fvad_batch_sync. */
int  fvad_batch_push_device(fvad_batch *b, const float *d_pcm, size_t stream_stride, size_t channel_stride,
                            size_t n_samples, uint64_t *first_index);
/* Wait for all submitted work; makes segments/events/traces readable. */
int  fvad_batch_sync(fvad_batch *b);
/* Back to the state right after create (same config, same device buffers). */
int  fvad_batch_reset(fvad_batch *b);

/* pipeline.vad.vad_machine.vad_segments (machine_idx 0) / alt_vad_machines[machine_idx-1]
 * (src/simulator/SimulationInstance.zig:211, VADMachine.zig:63).  The array is owned by the
 * handle and valid until the next push/reset/destroy.  Implies fvad_batch_sync. */
int  fvad_batch_segments(fvad_batch *b, size_t stream, int machine_idx, const fvad_segment **segs, size_t *n);
/* Recording-state results of the main machine, in order (drives the Recorder). */
int  fvad_batch_events(fvad_batch *b, size_t stream, const fvad_capture_event **events, size_t *n);
/* ---- Recorder: Callbacks.on_recording (src/AudioPipeline.zig:15-18,134-195; src/AudioPipeline/Recorder.zig:52-146) ----
 * Every started -> completed pair of the main machine yields one clip: the ORIGINAL samples
 * [start - 2 s (clamped at 0), end + 2 s) of the quietest channel (Recorder.findBestChannel: smallest f32 in-order RMS over
 * the clip, first channel wins ties), as the AudioBuffer the reference hands to on_recording
 * (Recorder.segmentToAudioBuffer, Recorder.zig:121-146).  An aborted recording yields nothing.
 * The reference copies the samples out of its ring while it records; here the decisions lag the input by the depth of
 * the tile pipeline, so the batch keeps a HISTORY of the original PCM in HBM (every push is copied into it on the copy
 * stream) and cuts the clips out of it with a gather kernel when the completing event has been seen and the stream has
 * reached the clip's end sample -- inside fvad_batch_sync (and wherever a push drains results).  history_samples = 0 keeps
 * everything pushed (the buffer doubles as needed: lossless, costs the input's size in HBM); otherwise a ring of that many
 * samples per channel: a clip whose start has been overwritten by then is delivered with lost = 1 and pcm = NULL (the
 * reference fails with IndexOutOfBounds when ITS 10-second ring is too short, MultiRingBuffer.zig:159-169).
 * Set before the first push.  on_recording (may be NULL) is called on the calling thread from inside push / sync, once per
 * clip in time order per stream; the clip is borrowed for the duration of the call (as AudioPipeline.zig:182-188 frees the
 * AudioBuffer right after the callback).  Without a callback, or in addition to it, fvad_batch_clips hands out the clips
 * completed since its previous call for that stream (owned by the handle until the next call for that stream). */
typedef struct fvad_clip {
    uint64_t global_start_frame_number;   /* AudioBuffer.global_start_frame_number = Recorder.startIndex */
    size_t   length;                      /* samples */
    size_t   sample_rate;
    size_t   channel;                     /* the channel findBestChannel chose (the AudioBuffer itself is mono) */
    float    duration_seconds;            /* (f32)length / (f32)sample_rate */
    int32_t  lost;                        /* 1: the history ring no longer held the start of the clip */
    const float *pcm;                     /* length samples of channel `channel`, or NULL when lost */
} fvad_clip;
typedef void (*fvad_on_recording_fn)(void *ctx, size_t stream, const fvad_clip *recording);
int  fvad_batch_set_recording(fvad_batch *b, size_t history_samples, fvad_on_recording_fn on_recording, void *ctx);
int  fvad_batch_clips(fvad_batch *b, size_t stream, const fvad_clip **clips, size_t *n);

/* Total samples written / consumed by the VAD so far (multi_ring_buffer.total_write_count,
 * VAD.pipeline_read_count). */
int  fvad_batch_counts(fvad_batch *b, uint64_t *total_write_count, uint64_t *pipeline_read_count);
/* The batch's main CUDA stream (a cudaStream_t): everything a push submits is ordered before work later
 * enqueued on it, so callers can bracket pushes with their own CUDA events (bench.py does). */
void *fvad_batch_cuda_stream(fvad_batch *b);

/* Optional traces for parity tests and for the config-5 sweep ("shared features").
 * Bits: 1 = per-STFT-frame records, 2 = denoised PCM, 4 = per-480-frame RNNoise VAD,
 * 8 = full STFT magnitude spectra.  Set before the first push. */
enum { FVAD_TRACE_RECORDS = 1, FVAD_TRACE_DENOISED = 2, FVAD_TRACE_RNN_VAD = 4, FVAD_TRACE_SPECTRA = 8,
       /* not a trace (no host copies, no synchronisation): the per-STFT-frame features of everything pushed stay in HBM
        * (12 B per frame per stream and 4 B per extra speech band) for fvad_batch_sweep */
       FVAD_KEEP_FEATURES = 16 };
int  fvad_batch_set_trace(fvad_batch *b, unsigned flags);
/* Copy out (host): returns the number of items available via *n; copies min(cap, *n). */
int  fvad_batch_trace_records(fvad_batch *b, size_t stream, fvad_stft_record *out, size_t cap, size_t *n);
int  fvad_batch_trace_denoised(fvad_batch *b, size_t stream, size_t channel, float *out, size_t cap, size_t *n);
int  fvad_batch_trace_rnn_vad(fvad_batch *b, size_t stream, float *out, size_t cap, size_t *n);
int  fvad_batch_trace_spectra(fvad_batch *b, size_t stream, size_t channel, float *out, size_t cap_floats, size_t *n_floats);

/* VAD.stateMachineStep's alt machines (src/AudioPipeline/VAD.zig:171-183,375-381) AFTER the fact and on the device:
 * n_configs further VADMachine configurations scanned over the features the batch kept of everything pushed so far
 * (FVAD_KEEP_FEATURES), one scan per (stream, config), no host round trip of the features (BASELINE configs[4]).  A
 * config's speech band must be one the batch was created with (main or alt config).  Output layout = fvad_vadm_sweep's:
 * segs[(stream*n_configs + k)*max_segs ...] in emission order, counts[stream*n_configs + k] (a count may exceed
 * max_segs: the list is truncated); host arrays, or device arrays when outputs_on_device != 0 (feed them to
 * fvad_eval_stats_batch_device).  Implies fvad_batch_sync; returns when the lists are complete. */
int  fvad_batch_sweep(fvad_batch *b, const fvad_vad_machine_config *configs, size_t n_configs, size_t max_segs,
                      fvad_segment *segs, uint32_t *counts, int outputs_on_device);
/* The kept features of one stream as records (host copy; the main machine's band): what FVAD_TRACE_RECORDS gives, without
 * its per-tile synchronisation.  Returns the number available via *n; copies min(cap, *n). */
int  fvad_batch_features(fvad_batch *b, size_t stream, fvad_stft_record *out, size_t cap, size_t *n);
/* STFT frames (feature records per stream) the machines have seen so far. */
int  fvad_batch_feature_count(fvad_batch *b, size_t *n_records);

/* ---- single pipeline = batch of 1 --------------------------------------------------- */
typedef struct fvad_batch fvad_pipeline;
int  fvad_pipeline_create(const fvad_config *cfg, fvad_pipeline **out);                 /* AudioPipeline.init */
int  fvad_pipeline_push(fvad_pipeline *p, const float *const *channel_pcm, size_t n_samples, uint64_t *first_index); /* pushSamples */
int  fvad_pipeline_segments(fvad_pipeline *p, int machine_idx, const fvad_segment **segs, size_t *n);
void fvad_pipeline_destroy(fvad_pipeline *p);                                           /* deinit */

/* ---- leaf drop-ins ------------------------------------------------------------------ */
/* FFT.zig: init/fft/binCount/binWidth/nyquistFreq/freqToBin/binToFreq/deinit (src/FFT.zig:28-144) */
typedef struct fvad_fft fvad_fft;
int    fvad_fft_create(size_t n_fft, size_t sample_rate, fvad_fft **out);               /* FFT.init; sizes as fvad_batch_create, up to 32768 */
void   fvad_fft_destroy(fvad_fft *f);                                                   /* FFT.deinit */
/* FFT.fft on a SplitSlice (first,second) of host memory; result = n_fft/2+1 normalised magnitudes. */
int    fvad_fft_run(fvad_fft *f, const float *first, size_t n_first, const float *second, size_t n_second,
                    const float *window, size_t n_window, float *result, size_t n_result);
size_t fvad_fft_bin_count(const fvad_fft *f);
float  fvad_fft_bin_width(const fvad_fft *f);
float  fvad_fft_nyquist_freq(const fvad_fft *f);
int    fvad_fft_freq_to_bin(const fvad_fft *f, float freq, size_t *bin);
int    fvad_fft_bin_to_freq(const fvad_fft *f, size_t bin, float *freq);
/* window_fn.hannWindowPeriodic / windowNormFactor (src/audio_utils/window_fn.zig:8-28,51-68), host */
void   fvad_hann_window_periodic(float *w, size_t n);
float  fvad_window_norm_factor(const float *w, size_t n);

/* Batched STFT (PipelineFFT.fft + averageVolumeInBand, src/AudioPipeline/PipelineFFT.zig:88-112)
 * over DEVICE memory: frame f of row r starts at d_pcm[r*row_stride + f*hop]; n_fft samples each;
 * periodic Hann window.  d_band_sum[r*n_frames+f] = sum of bins min_bin..max_bin (inclusive) of the
 * normalised magnitudes; d_spectra (may be NULL) gets [r][f][n_fft/2+1] magnitudes. */
int    fvad_stft_batch(fvad_fft *f, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames,
                       size_t hop, size_t min_bin, size_t max_bin, float *d_band_sum, float *d_spectra,
                       void *cuda_stream);

/* Spectrogram.compute (src/audio_utils/Spectrogram.zig:30-94): the second consumer of FFT.fft.
 * width = (n_samples - hop_size) / hop_size frames (incomplete frames dropped, :39), frame i covers
 * samples[i*hop_size, i*hop_size + fft_size), periodic Hann window (the reference ignores
 * SpectrogramOptions.window_function, :49-50), values[i*height + k] = normalised magnitude of bin k.
 * `samples` and `values` are HOST arrays (the reference's []f32 in, []f32 out); `bin_labels` (height floats,
 * FFT.binToFreq) may be NULL.  fvad_spectrogram_shape does no compute and needs no device.
 * Errors: InsufficientSamples (n_samples < fft_size, :34-36); OutOfRange where the reference's slice
 * samples[from..to] would run past the input (hop_size < fft_size/2: a bounds panic in the reference);
 * Unsupported for an fft_size whose half has a prime factor above 5, or an odd hop_size with fft_size 2048. */
typedef struct fvad_spectrogram_info {
    size_t height;          /* bins = fft_size/2 + 1 */
    size_t width;           /* steps */
    float  bin_width;
    float  nyquist_freq;
    double length_sec;      /* processed samples / sample_rate (:81) */
} fvad_spectrogram_info;
int    fvad_spectrogram_shape(size_t n_samples, size_t sample_rate, size_t fft_size, size_t hop_size,
                              fvad_spectrogram_info *info);
int    fvad_spectrogram_compute(const float *samples, size_t n_samples, size_t sample_rate, size_t fft_size,
                                size_t hop_size, float *values, size_t values_len, float *bin_labels,
                                fvad_spectrogram_info *info);

/* Denoiser.zig: init/denoise/getFrameSize/deinit (src/Denoiser.zig:22-70) */
typedef struct fvad_denoiser fvad_denoiser;
int    fvad_denoiser_create(fvad_denoiser **out);
void   fvad_denoiser_destroy(fvad_denoiser *d);
size_t fvad_denoiser_frame_size(void);                                                   /* 480 */
/* Denoiser.denoise on a SplitSlice; n_first+n_second must be 480, n_result 480. Host memory. */
int    fvad_denoiser_denoise(fvad_denoiser *d, const float *first, size_t n_first, const float *second,
                             size_t n_second, float *result, size_t n_result, float *vad);
/* Many independent denoisers x many consecutive frames, HOST memory:
 * pcm[s] -> n_frames*480 normalised samples; out[s] likewise; vad[s] -> n_frames. */
typedef struct fvad_denoiser_batch fvad_denoiser_batch;
int    fvad_denoiser_batch_create(size_t n_streams, int device, fvad_denoiser_batch **out);
void   fvad_denoiser_batch_destroy(fvad_denoiser_batch *d);
int    fvad_denoiser_batch_run(fvad_denoiser_batch *d, const float *const *pcm, size_t n_frames,
                               float *const *out, float *const *vad);

/* ---- VADMachine sweep over shared features (BASELINE config 5; VAD.zig:171-183,375-380) - */
/* records: host [n_streams][n_records]; configs: host [n_configs].  Output: for (stream s,
 * config k) the segments land at segs[(s*n_configs+k)*max_segs ...] and their count at
 * counts[s*n_configs+k] (count may exceed max_segs: truncated). */
int    fvad_vadm_sweep(const fvad_stft_record *records, size_t n_streams, size_t n_records,
                       const fvad_vad_machine_config *configs, size_t n_configs, size_t sample_rate,
                       size_t fft_size, int device, fvad_segment *segs, size_t max_segs, uint32_t *counts);

/* The same sweep for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Calls no cudaMalloc / cudaFree / allocating library routine: the scans' scratch (their f64 rings) is one
 * file-scope device array of FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES that every GPU loading the library holds
 * until exit.  Runs on the CURRENT device, asynchronously on `cuda_stream` (which must belong to it).
 * Inputs are DEVICE planes [n_streams][n_records] of the records' fields (min_volume, vol_ratio, rnn_vad;
 * d_rnn_vad may be NULL = fft_input.vad == null); record f has fft_result.index = first_index + f*fft_size.
 * Outputs are DEVICE arrays in fvad_vadm_sweep's layout: d_segs[(s*n_configs+k)*max_segs ...], d_counts[s*n_configs+k]
 * (a count may exceed max_segs: truncated).
 * Limits: n_configs <= FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS and
 *   n_streams*n_configs * (max long + max short + max ratio ring length over the configs) * 8 B
 *     <= FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES - 32 KiB        (default config: 4233 entries -> 494 scans);
 * beyond them the call fails with FVAD_EOUT_OF_MEMORY and a message naming the limit.
 * All calls share the one array, which keeps the last call's contents (the kernel rewrites what it uses first):
 * calls that overlap on the GPU race -- serialise them on one stream or with events. */
#define FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES (16u << 20)
#define FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS   256
int    fvad_vadm_sweep_static(const float *d_min_volume, const float *d_vol_ratio, const float *d_rnn_vad,
                              size_t n_streams, size_t n_records, uint64_t first_index,
                              const fvad_vad_machine_config *configs, size_t n_configs, size_t sample_rate,
                              size_t fft_size, fvad_segment *d_segs, size_t max_segs, uint32_t *d_counts,
                              void *cuda_stream);

/* ---- Evaluator statistics over many result sets (the sweep's consumer) ------------
 * statistics.fromEvaluator (src/Evaluator/statistics.zig:85-113) on top of Evaluator.initAndRun's overlap matching
 * (src/Evaluator.zig:90-156; SpeechSegment.findOverlapping/overlapWith, src/Evaluator/SpeechSegment.zig:22-49) with
 * the extrusion / gap filling of extrudeSegments (statistics.zig:225-250), all in f32 in the reference's order.
 * One device thread per result set.  Set i has counts[i] VAD segments at segs[i*max_segs ...] (the layout
 * fvad_vadm_sweep writes; seconds = (f32)sample / (f32)sample_rate as SimulationInstance.storeResult does,
 * src/simulator/SimulationInstance.zig:226-227) and is scored against reference list set_ref[i]: labels
 * ref_from/ref_to[ref_offsets[j] .. ref_offsets[j+1]) in seconds.  Both sides must be ordered by start
 * (initAndRun sorts them; VADMachine emits in time order, Audacity labels are sorted by fvad's host code). */
typedef struct fvad_stat_config {           /* statistics.StatConfig, statistics.zig:77-83 */
    float ignore_shorter_than_sec;
    float extrude_start;
    float extrude_end;
    float fill_gaps;
} fvad_stat_config;
typedef struct fvad_single_stats {          /* statistics.SingleStats, statistics.zig:8-39 */
    float total_positives_sec, true_positives_sec, false_positives_sec, false_negatives_sec;
    float true_positive_rate, false_negative_rate, false_discovery_rate, precision;
    float fm_index, f_score, f_score_beta;
} fvad_single_stats;
int    fvad_eval_stats_batch(const fvad_segment *segs, size_t n_sets, size_t max_segs, const uint32_t *counts,
                             const uint32_t *set_ref, const float *ref_from, const float *ref_to,
                             const uint32_t *ref_offsets, size_t n_refs, size_t sample_rate,
                             const fvad_stat_config *config, int device, fvad_single_stats *out);
/* The same with the result sets already on the device (d_segs, d_counts: what fvad_batch_sweep / fvad_vadm_sweep_static
 * leave there); everything else is host memory as above.  Runs on the current device. */
int    fvad_eval_stats_batch_device(const fvad_segment *d_segs, size_t n_sets, size_t max_segs, const uint32_t *d_counts,
                                    const uint32_t *set_ref, const float *ref_from, const float *ref_to,
                                    const uint32_t *ref_offsets, size_t n_refs, size_t sample_rate,
                                    const fvad_stat_config *config, fvad_single_stats *out);

/* ---- seeded synthetic 48 kHz input ------------------------------ */
/* Bit-identical on host and device.  Sample i of stream `stream_id` (absolute index
 * first_sample + i).  kind 0 = "onboard": engine bed + radio-speech bursts + gaps. */
int    fvad_synth_fill_host(float *dst, uint64_t seed, uint64_t stream_id, uint64_t first_sample, size_t n);
int    fvad_synth_fill_device(float *d_dst, size_t stream_stride, uint64_t seed, uint64_t first_stream_id,
                              size_t n_streams, uint64_t first_sample, size_t n, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* FVAD_H */
