/*
 * oracle.h -- CPU parity oracle for the Formula-VAD AudioPipeline hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under formula-vad_b200/ (the product) may
 * include, link or call this.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Two builds of the same interface (see oracle/Makefile):
 *   liboracle_port.so      "port": plain-C restatement of the reference's arithmetic
 *                          (KissFFT real FFT, RNNoise, and the Zig glue).  Builds anywhere.
 *   _ref/liboracle_ref.so  "reference": the SAME Zig-glue restatement, but the leaf
 *                          arithmetic is the reference's own C sources compiled
 *                          unmodified from /root/reference (lib/kissfft, lib/rnnoise).
 *                          The Zig layer cannot be built (no zig toolchain), so the
 *                          glue is always the restatement.
 *
 * Parity pinning: the reference ships no golden vectors for this path
 * (SURVEY.md section 4 / 8c).  The port is pinned bit-for-bit against the
 * reference build on seeded synthetic input (tests/test_oracle_*.py) and
 * against committed fixtures under tests/golden/ generated from the
 * reference build (tools/make_golden.py).
 */
#ifndef FVAD_ORACLE_H
#define FVAD_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Mirrors VADMachine.Config (src/AudioPipeline/VADMachine.zig:18-39). */
typedef struct {
    float  speech_min_freq;
    float  speech_max_freq;
    float  long_term_speech_avg_sec;
    int    has_initial_long_term_avg;   /* ?f64: 0 = null */
    double initial_long_term_avg;
    float  short_term_speech_avg_sec;
    float  speech_threshold_factor;
    float  channel_vol_ratio_avg_sec;
    float  channel_vol_ratio_threshold;
    float  min_consecutive_sec_to_open;
    float  max_speech_gap_sec;
    float  min_vad_duration_sec;
} orc_vadm_config;

/* Mirrors VAD.VADSpeechSegment (src/AudioPipeline/VAD.zig:25-30). */
typedef struct {
    uint64_t sample_from;
    uint64_t sample_to;
    float    debug_rnn_vad;
    float    debug_avg_speech_vol_ratio;
} orc_segment;

/* Per STFT-frame record handed to the state machines (what VADMachine.run sees). */
typedef struct {
    uint64_t index;        /* fft_result.index */
    float    min_volume;   /* min over channels of the speech-band sum (main machine's band) */
    float    vol_ratio;    /* fft_input.volume_ratio */
    float    rnn_vad;      /* fft_input.vad orelse 0 */
    float    has_vad;      /* 1 if fft_input.vad != null */
} orc_stft_record;

void orc_default_vadm_config(orc_vadm_config *c);

/* ---- leaf: FFT.zig ------------------------------------------------------ */
/* FFT.fft (src/FFT.zig:70-98) on a split slice.  faithful!=0 regenerates the
 * KissFFT state on every call as FFT.zig:84 does.  Returns 0 or a negative
 * error (-1 InvalidSamplesLength, -2 InvalidWindowLength, -3 InvalidResultLength,
 * -4 InvalidFFTSize). */
typedef struct orc_fft orc_fft;
orc_fft *orc_fft_create(int n_fft, int sample_rate, int faithful);
void     orc_fft_destroy(orc_fft *f);
int      orc_fft_run(orc_fft *f, const float *first, size_t n_first, const float *second,
                     size_t n_second, const float *window, size_t n_window, float *result,
                     size_t n_result);
/* raw complex spectrum of the last orc_fft_run (n_fft/2+1 interleaved re,im) */
const float *orc_fft_last_spectrum(const orc_fft *f);
long     orc_fft_freq_to_bin(const orc_fft *f, float freq);   /* <0 = error */
void     orc_hann_window_periodic(float *w, size_t n);        /* window_fn.zig:22-28,51-68 */
float    orc_window_norm_factor(const float *w, size_t n);    /* window_fn.zig:8-16 */
float    orc_rms_volume(const float *first, size_t n_first, const float *second, size_t n_second);

/* ---- leaf: Denoiser.zig -------------------------------------------------- */
typedef struct orc_denoiser orc_denoiser;
orc_denoiser *orc_denoiser_create(void);
void          orc_denoiser_destroy(orc_denoiser *d);
/* Denoiser.denoise (src/Denoiser.zig:45-66): in/out normalised [-1,1] PCM, 480 samples. */
float         orc_denoiser_denoise(orc_denoiser *d, const float *in480, float *out480);

/* Per-frame trace of the RNNoise internals (port build only; the reference
 * build fills what it can reach through the library's non-static functions). */
typedef struct {
    float hp[480];            /* biquad output (denoise.c:470-472) */
    float X[2 * 481];         /* analysis spectrum before gains */
    float Ex[22];
    float Ep[22];
    float Exp[22];
    int   pitch_index;
    float pitch_gain;
    int   silence;
    float features[42];
    float gains[22];          /* RNN output g before the lastg floor */
    float vad_prob;
    float out[480];           /* rnnoise_process_frame output (s16 scale) */
} orc_rnnoise_trace;
/* rnnoise_process_frame on s16-scaled floats, optional trace (may be NULL). */
float orc_denoiser_process_raw(orc_denoiser *d, const float *in480, float *out480,
                               orc_rnnoise_trace *trace);

/* ---- the pipeline -------------------------------------------------------- */
typedef struct orc_pipeline orc_pipeline;
/* AudioPipeline.init (src/AudioPipeline.zig:39-75) + VAD.init (VAD.zig:97-192).
 * Returns NULL and sets *err (-10 InvalidSampleRate, -4 InvalidFFTSize). */
orc_pipeline *orc_pipeline_create(int sample_rate, int n_channels, int fft_size, int use_denoiser,
                                  const orc_vadm_config *main_cfg, const orc_vadm_config *alt_cfgs,
                                  int n_alt, int faithful, int *err);
void     orc_pipeline_destroy(orc_pipeline *p);
/* AudioPipeline.pushSamples (src/AudioPipeline.zig:86-120); returns first sample index. */
uint64_t orc_pipeline_push(orc_pipeline *p, const float *const *channel_pcm, size_t n);
/* machine_idx 0 = vad_machine, 1.. = alt_vad_machines.  Copies up to cap segments; returns count. */
size_t   orc_pipeline_segments(const orc_pipeline *p, int machine_idx, orc_segment *out, size_t cap);
/* Recording-state events of machine 0 (VAD.zig:362-373): kind 1 started, 2 completed, 3 aborted. */
typedef struct { uint64_t at_index; int kind; uint64_t sample_number; } orc_capture_event;
size_t   orc_pipeline_events(const orc_pipeline *p, orc_capture_event *out, size_t cap);
/* Tracing: keep per-STFT-frame records / denoised PCM (channel-planar) for comparison. */
void     orc_pipeline_enable_trace(orc_pipeline *p, int stft_records, int denoised_pcm, int rnn_vad);
size_t   orc_pipeline_stft_records(const orc_pipeline *p, orc_stft_record *out, size_t cap);
size_t   orc_pipeline_denoised(const orc_pipeline *p, int channel, float *out, size_t cap);
size_t   orc_pipeline_rnn_vad(const orc_pipeline *p, float *out, size_t cap);
/* band sums of the last STFT frame's spectrum and the full spectra trace */
size_t   orc_pipeline_spectra(const orc_pipeline *p, int channel, float *out, size_t cap_floats);
void     orc_pipeline_enable_spectra(orc_pipeline *p, int on);

/* VADMachine alone over precomputed records (config-5 sweep checker). */
size_t   orc_vadm_scan(const orc_vadm_config *cfg, int sample_rate, int fft_size,
                       const orc_stft_record *rec, size_t n_rec, orc_segment *out, size_t cap);

const char *orc_build_kind(void);   /* "port" or "reference" */

#ifdef __cplusplus
}
#endif
#endif
