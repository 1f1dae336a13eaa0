/*
 * glue.c -- plain-C restatement of the reference's Zig layer for the hot path:
 *   src/FFT.zig, src/Denoiser.zig, src/audio_utils.zig, src/audio_utils/window_fn.zig,
 *   src/structures/RollingAverage.zig, src/AudioPipeline/{VAD,VADMachine,PipelineFFT,
 *   SegmentWriter}.zig, src/AudioPipeline.zig (pushSamples).
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Compiled with -ffp-contract=off.
 * The Zig toolchain is absent on this box, so this layer is a restatement in both
 * oracle builds; the leaf arithmetic below it is either the port or the reference.
 */
#include "leaf.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_default_vadm_config(orc_vadm_config *c) {   /* VADMachine.zig:18-39 */
    c->speech_min_freq = 100; c->speech_max_freq = 1500;
    c->long_term_speech_avg_sec = 180;
    c->has_initial_long_term_avg = 1; c->initial_long_term_avg = 0.005;
    c->short_term_speech_avg_sec = 0.2f;
    c->speech_threshold_factor = 18;
    c->channel_vol_ratio_avg_sec = 0.5f; c->channel_vol_ratio_threshold = 0.5f;
    c->min_consecutive_sec_to_open = 0.2f;
    c->max_speech_gap_sec = 2;
    c->min_vad_duration_sec = 0.7f;
}

/* ---- window_fn.zig -------------------------------------------------------- */
/* hannWindowPeriodic -> cosineSumWindowPeriodic (window_fn.zig:22-28,51-68), K=1,
 * alphas {0.5,0.5}, all f32.  `@cos` on f32 is Zig's compiler-rt cosf, a port of musl's:
 * restated in musl_cosf.c (one entry of the 2048-entry table differs from the correctly
 * rounded cosine; tests/test_oracle.py::test_hann_window_cosf_pin). */
float orc_musl_cosf(float x);
void orc_hann_window_periodic(float *w, size_t n) {
    const float N = (float)n;
    const float two_pi = (float)(2 * 3.14159265358979323846264338327950288419716939937510);
    for (size_t i = 0; i < n; i++) {
        float nf = (float)i;
        float r = 0;
        /* k = 0: pow(-1,0) * a0 * cos(0) */
        r += (1.0f * 0.5f) * orc_musl_cosf(((two_pi * 0.0f) * nf) / N);
        /* k = 1: pow(-1,1) * a1 * cos(2 pi n / N) */
        r += (-1.0f * 0.5f) * orc_musl_cosf(((two_pi * 1.0f) * nf) / N);
        w[i] = r;
    }
}
float orc_window_norm_factor(const float *w, size_t n) {   /* window_fn.zig:8-16 */
    float sum = 0;
    for (size_t i = 0; i < n; i++) sum += w[i];
    return (float)n / sum;
}
float orc_rms_volume(const float *first, size_t n_first, const float *second, size_t n_second) { /* audio_utils.zig:14-24 */
    float sum = 0;
    for (size_t i = 0; i < n_first; i++) sum += first[i] * first[i];
    for (size_t i = 0; i < n_second; i++) sum += second[i] * second[i];
    float mean = sum / (float)(n_first + n_second);
    return sqrtf(mean);
}

/* ---- FFT.zig ---------------------------------------------------------------- */
struct orc_fft { int n_fft, sample_rate, faithful; leaf_rfft *leaf; float *f_in, *f_out; };
orc_fft *orc_fft_create(int n_fft, int sample_rate, int faithful) {
    if (n_fft <= 0 || (n_fft % 2) != 0) return NULL;           /* FFT.zig:29-31 */
    leaf_rfft *leaf = leaf_rfft_create(n_fft);
    if (!leaf) return NULL;
    orc_fft *f = (orc_fft *)calloc(1, sizeof(*f));
    f->n_fft = n_fft; f->sample_rate = sample_rate; f->faithful = faithful; f->leaf = leaf;
    f->f_in = (float *)malloc(sizeof(float) * n_fft);
    f->f_out = (float *)malloc(sizeof(float) * 2 * n_fft);
    return f;
}
void orc_fft_destroy(orc_fft *f) { if (f) { leaf_rfft_destroy(f->leaf); free(f->f_in); free(f->f_out); free(f); } }
const float *orc_fft_last_spectrum(const orc_fft *f) { return f->f_out; }
int orc_fft_run(orc_fft *f, const float *first, size_t n_first, const float *second, size_t n_second,
                const float *window, size_t n_window, float *result, size_t n_result) {
    size_t n = (size_t)f->n_fft, bins = n / 2 + 1;
    if (n_first + n_second != n) return -1;
    if (n_window != n) return -2;
    if (n_result != bins) return -3;
    if (f->faithful) leaf_rfft_reset(f->leaf);                   /* FFT.zig:84 */
    for (size_t i = 0; i < n_first; i++) f->f_in[i] = first[i] * window[i];             /* FFT.zig:147-159 */
    for (size_t i = 0; i < n_second; i++) f->f_in[n_first + i] = second[i] * window[n_first + i];
    leaf_rfft_run(f->leaf, f->f_in, f->f_out);                   /* FFT.zig:90 */
    float window_norm = orc_window_norm_factor(window, n);       /* FFT.zig:94 */
    float norm_factor = window_norm / (float)(n / 2);            /* FFT.zig:165 */
    for (size_t i = 0; i < bins; i++) {                          /* FFT.zig:167-176 */
        float r = f->f_out[2 * i], im = f->f_out[2 * i + 1];
        float r2 = r * r, i2 = im * im;
        result[i] = sqrtf(r2 + i2) * norm_factor;
    }
    return 0;
}
long orc_fft_freq_to_bin(const orc_fft *f, float freq) {        /* FFT.zig:120-131 */
    float nyq = (float)f->sample_rate / 2;
    if (freq > nyq) return -5;
    if (freq < 0) return -6;
    float bw = (float)f->sample_rate / (float)f->n_fft;
    return (long)roundf(freq / bw);
}

/* ---- Denoiser.zig ------------------------------------------------------------ */
struct orc_denoiser { leaf_rnn *leaf; float in[480]; };
orc_denoiser *orc_denoiser_create(void) {
    orc_denoiser *d = (orc_denoiser *)calloc(1, sizeof(*d));
    d->leaf = leaf_rnn_create();
    return d;
}
void orc_denoiser_destroy(orc_denoiser *d) { if (d) { leaf_rnn_destroy(d->leaf); free(d); } }
float orc_denoiser_denoise(orc_denoiser *d, const float *in480, float *out480) {   /* Denoiser.zig:45-94 */
    const float scalar = 32767.0f, inv = 1.0f / 32767.0f;
    for (int i = 0; i < 480; i++) d->in[i] = in480[i] * scalar;
    float vad = leaf_rnn_process(d->leaf, out480, d->in, NULL);
    for (int i = 0; i < 480; i++) out480[i] *= inv;
    return vad;
}
float orc_denoiser_process_raw(orc_denoiser *d, const float *in480, float *out480, orc_rnnoise_trace *trace) {
    return leaf_rnn_process(d->leaf, out480, in480, trace);
}

/* ---- RollingAverage.zig ------------------------------------------------------- */
typedef struct { double *data; size_t len, write_idx, written; int has_last; double last_avg; } rollavg;
static double ra_avg(rollavg *r) {                               /* RollingAverage.zig:45-56 */
    double a = 0.0, scalar = 1.0 / (double)r->written;
    for (size_t i = 0; i < r->written; i++) a += r->data[i] * scalar;
    r->last_avg = a; r->has_last = 1;
    return a;
}
static void ra_init(rollavg *r, size_t count, int has_init, double init) {   /* RollingAverage.zig:11-28 */
    memset(r, 0, sizeof(*r));
    r->data = (double *)calloc(count ? count : 1, sizeof(double));
    r->len = count;
    if (has_init) {
        for (size_t i = 0; i < count; i++) r->data[i] = init;
        r->written = count;
        ra_avg(r);
    }
}
static double ra_push(rollavg *r, float sample) {                /* RollingAverage.zig:34-43 */
    r->data[r->write_idx] = sample;
    r->write_idx = (r->write_idx + 1) % r->len;
    if (r->written < r->len) r->written++;
    return ra_avg(r);
}

/* ---- VADMachine.zig ------------------------------------------------------------ */
enum { ST_CLOSED, ST_OPENING, ST_OPEN, ST_CLOSING };
typedef struct {
    orc_vadm_config cfg;
    int sample_rate;
    long min_bin, max_bin;
    int state;
    rollavg long_term, short_term, vol_ratio;
    uint64_t speech_start_index, speech_end_index;
    float speech_rnn_vad; size_t speech_rnn_vad_count;
    float speech_vol_ratio; size_t speech_vol_ratio_count;
    orc_segment *segs; size_t n_segs, cap_segs;
} vadm;

static int vadm_init(vadm *m, const orc_vadm_config *cfg, int sample_rate, int fft_size) {   /* VADMachine.zig:65-117 */
    memset(m, 0, sizeof(*m));
    m->cfg = *cfg; m->sample_rate = sample_rate;
    float eval_per_sec = (float)sample_rate / (float)fft_size;
    size_t long_len = (size_t)(eval_per_sec * cfg->long_term_speech_avg_sec);
    size_t short_len = (size_t)(eval_per_sec * cfg->short_term_speech_avg_sec);
    size_t ratio_len = (size_t)(eval_per_sec * cfg->channel_vol_ratio_avg_sec);
    if (ratio_len == 0) return -7;              /* the reference would index an empty ring */
    ra_init(&m->long_term, long_len > 1 ? long_len : 1, cfg->has_initial_long_term_avg, cfg->initial_long_term_avg);
    ra_init(&m->short_term, short_len > 1 ? short_len : 1, 0, 0);
    ra_init(&m->vol_ratio, ratio_len, 0, 0);
    m->state = ST_CLOSED;
    return 0;
}
static void vadm_free(vadm *m) { free(m->long_term.data); free(m->short_term.data); free(m->vol_ratio.data); free(m->segs); }

static uint64_t off_start(const vadm *m, uint64_t from) {         /* VADMachine.zig:297-302 */
    uint64_t buf = (uint64_t)((float)m->sample_rate * 2);
    return buf > from ? 0 : from - buf;
}
static uint64_t off_end(const vadm *m, uint64_t to) {             /* VADMachine.zig:305-310 */
    return to + (uint64_t)((float)m->sample_rate * 2);
}
static void track_stats(vadm *m, float vad, float ratio, int from, int to) {   /* VADMachine.zig:232-250 */
    if (from == ST_CLOSED && to == ST_OPENING) {
        m->speech_rnn_vad = vad; m->speech_rnn_vad_count = 1;
        m->speech_vol_ratio = ratio; m->speech_vol_ratio_count = 1;
    } else if (from == ST_OPENING || from == ST_OPEN) {
        m->speech_rnn_vad += vad; m->speech_rnn_vad_count += 1;
        m->speech_vol_ratio += ratio; m->speech_vol_ratio_count += 1;
    }
}
/* One VADMachine.run (VADMachine.zig:126-230) given the already reduced speech-band
 * minimum.  Returns the recording state (0 none, 1 started, 2 completed, 3 aborted). */
static int vadm_step(vadm *m, float min_volume, float volume_ratio, float vad_or_0, uint64_t index, uint64_t *sample_number) {
    const orc_vadm_config *c = &m->cfg;
    float sr = (float)m->sample_rate;
    uint64_t min_consecutive = (uint64_t)(sr * c->min_consecutive_sec_to_open);
    uint64_t max_gap = (uint64_t)(sr * c->max_speech_gap_sec);
    double short_term = ra_push(&m->short_term, min_volume);
    double cvr = ra_push(&m->vol_ratio, volume_ratio);
    double base = m->long_term.has_last ? m->long_term.last_avg
                : (c->has_initial_long_term_avg ? c->initial_long_term_avg : short_term);
    double threshold = base * (double)c->speech_threshold_factor;
    int met = short_term > threshold && cvr > (double)c->channel_vol_ratio_threshold;
    if (!met) ra_push(&m->long_term, min_volume);
    int rec = 0; *sample_number = 0;
    switch (m->state) {
    case ST_CLOSED:
        if (met) { m->state = ST_OPENING; m->speech_start_index = index; }
        track_stats(m, vad_or_0, volume_ratio, ST_CLOSED, m->state);
        break;
    case ST_OPENING: {
        uint64_t since = index - m->speech_start_index;
        if (met && since >= min_consecutive) { m->state = ST_OPEN; rec = 1; *sample_number = off_start(m, m->speech_start_index); }
        else if (!met) m->state = ST_CLOSED;
        track_stats(m, vad_or_0, volume_ratio, ST_OPENING, m->state);
        break; }
    case ST_OPEN:
        if (!met) { m->state = ST_CLOSING; m->speech_end_index = index; }
        track_stats(m, vad_or_0, volume_ratio, ST_OPEN, m->state);
        break;
    case ST_CLOSING: {
        uint64_t since = index - m->speech_end_index;
        if (met) m->state = ST_OPEN;
        else if (since >= max_gap) {
            m->state = ST_CLOSED;
            /* onSpeechEnd, VADMachine.zig:252-294 */
            uint64_t from = m->speech_start_index, to = m->speech_end_index;
            float length_realtime = (float)(to - from) / sr;
            int dur_met = length_realtime >= c->min_vad_duration_sec;
            float avg_vad = m->speech_rnn_vad / (float)m->speech_rnn_vad_count;
            float avg_ratio = m->speech_vol_ratio / (float)m->speech_vol_ratio_count;
            if (dur_met) {
                if (m->n_segs == m->cap_segs) {
                    m->cap_segs = m->cap_segs ? m->cap_segs * 2 : 128;
                    m->segs = (orc_segment *)realloc(m->segs, m->cap_segs * sizeof(orc_segment));
                }
                orc_segment s = { off_start(m, from), off_end(m, to), avg_vad, avg_ratio };
                m->segs[m->n_segs++] = s;
                rec = 2; *sample_number = off_end(m, to);
            } else { rec = 3; *sample_number = 0; }
        }
        track_stats(m, vad_or_0, volume_ratio, ST_CLOSING, m->state);
        break; }
    }
    return rec;
}

size_t orc_vadm_scan(const orc_vadm_config *cfg, int sample_rate, int fft_size, const orc_stft_record *rec,
                     size_t n_rec, orc_segment *out, size_t cap) {
    vadm m;
    if (vadm_init(&m, cfg, sample_rate, fft_size)) return 0;
    for (size_t i = 0; i < n_rec; i++) {
        uint64_t sn;
        vadm_step(&m, rec[i].min_volume, rec[i].vol_ratio, rec[i].has_vad != 0 ? rec[i].rnn_vad : 0.0f, rec[i].index, &sn);
    }
    size_t n = m.n_segs;
    for (size_t i = 0; i < n && i < cap; i++) out[i] = m.segs[i];
    vadm_free(&m);
    return n;
}

/* ---- AudioPipeline.zig + VAD.zig ------------------------------------------------ */
typedef struct { float *v; size_t n, cap; } fvec;
static void fvec_push(fvec *a, const float *x, size_t n) {
    if (a->n + n > a->cap) { a->cap = (a->n + n) * 2 + 1024; a->v = (float *)realloc(a->v, a->cap * sizeof(float)); }
    memcpy(a->v + a->n, x, n * sizeof(float)); a->n += n;
}

struct orc_pipeline {
    int sample_rate, n_channels, fft_size, use_denoiser, faithful;
    uint64_t total_write_count, read_count;          /* multi_ring_buffer.total_write_count, VAD.pipeline_read_count */
    fvec *pending; size_t pending_off;                /* stands in for the ring: samples not yet consumed */
    orc_denoiser *denoiser;                           /* ONE state shared by all channels (VAD.zig:75,108) */
    float **denoised;                                 /* temp_denoiser_segment [C][480] */
    float **fft_buf; size_t fft_write; uint64_t fft_index;   /* fft_input_buffer (SegmentWriter) */
    float tmp_rnn_vad, tmp_vol_ratio;
    orc_fft *fft; float *window; float **bins;        /* pipeline_fft + temp_pipeline_fft_result */
    int n_machines; vadm *machines;
    orc_capture_event *events; size_t n_events, cap_events;
    int tr_rec, tr_pcm, tr_vad, tr_spec;
    orc_stft_record *recs; size_t n_recs, cap_recs;
    fvec *pcm_trace; fvec vad_trace; fvec *spec_trace;
};

orc_pipeline *orc_pipeline_create(int sample_rate, int n_channels, int fft_size, int use_denoiser,
                                  const orc_vadm_config *main_cfg, const orc_vadm_config *alt_cfgs, int n_alt,
                                  int faithful, int *err) {
    int e_dummy; if (!err) err = &e_dummy; *err = 0;
    if (sample_rate != 48000) { *err = -10; return NULL; }        /* VAD.zig:101-104 */
    orc_fft *fft = orc_fft_create(fft_size, sample_rate, faithful);
    if (!fft) { *err = -4; return NULL; }
    orc_pipeline *p = (orc_pipeline *)calloc(1, sizeof(*p));
    p->sample_rate = sample_rate; p->n_channels = n_channels; p->fft_size = fft_size;
    p->use_denoiser = use_denoiser; p->faithful = faithful; p->fft = fft;
    p->pending = (fvec *)calloc(n_channels, sizeof(fvec));
    p->denoiser = orc_denoiser_create();                          /* VAD.zig:108 (created even if unused) */
    p->denoised = (float **)calloc(n_channels, sizeof(float *));
    p->fft_buf = (float **)calloc(n_channels, sizeof(float *));
    p->bins = (float **)calloc(n_channels, sizeof(float *));
    p->pcm_trace = (fvec *)calloc(n_channels, sizeof(fvec));
    p->spec_trace = (fvec *)calloc(n_channels, sizeof(fvec));
    for (int c = 0; c < n_channels; c++) {
        p->denoised[c] = (float *)calloc(480, sizeof(float));
        p->fft_buf[c] = (float *)calloc(fft_size, sizeof(float));
        p->bins[c] = (float *)calloc(fft_size / 2 + 1, sizeof(float));
    }
    p->window = (float *)malloc(sizeof(float) * fft_size);
    orc_hann_window_periodic(p->window, fft_size);                /* PipelineFFT.zig:66-68 */
    p->n_machines = 1 + n_alt;
    p->machines = (vadm *)calloc(p->n_machines, sizeof(vadm));
    for (int k = 0; k < p->n_machines; k++) {
        const orc_vadm_config *c = k == 0 ? main_cfg : &alt_cfgs[k - 1];
        int e = vadm_init(&p->machines[k], c, sample_rate, fft_size);
        long lo = orc_fft_freq_to_bin(fft, c->speech_min_freq), hi = orc_fft_freq_to_bin(fft, c->speech_max_freq);
        if (e || lo < 0 || hi < 0) { *err = e ? e : (int)(lo < 0 ? lo : hi); orc_pipeline_destroy(p); return NULL; }
        p->machines[k].min_bin = lo; p->machines[k].max_bin = hi;
    }
    return p;
}

void orc_pipeline_destroy(orc_pipeline *p) {
    if (!p) return;
    for (int c = 0; c < p->n_channels; c++) {
        free(p->pending[c].v); free(p->denoised[c]); free(p->fft_buf[c]); free(p->bins[c]);
        free(p->pcm_trace[c].v); free(p->spec_trace[c].v);
    }
    for (int k = 0; k < p->n_machines; k++) vadm_free(&p->machines[k]);
    free(p->pending); free(p->denoised); free(p->fft_buf); free(p->bins); free(p->pcm_trace); free(p->spec_trace);
    free(p->machines); free(p->events); free(p->recs); free(p->vad_trace.v); free(p->window);
    orc_denoiser_destroy(p->denoiser); orc_fft_destroy(p->fft); free(p);
}

void orc_pipeline_enable_trace(orc_pipeline *p, int stft_records, int denoised_pcm, int rnn_vad) {
    p->tr_rec = stft_records; p->tr_pcm = denoised_pcm; p->tr_vad = rnn_vad;
}
void orc_pipeline_enable_spectra(orc_pipeline *p, int on) { p->tr_spec = on; }

/* fftStep + stateMachineStep (VAD.zig:350-381): segment = C x fft_size samples */
static void fft_step(orc_pipeline *p, float *const *seg, uint64_t index, int has_vad, float vad, float volume_ratio) {
    size_t bins = (size_t)p->fft_size / 2 + 1;
    for (int c = 0; c < p->n_channels; c++) {                      /* PipelineFFT.zig:88-97 */
        orc_fft_run(p->fft, seg[c], (size_t)p->fft_size, NULL, 0, p->window, (size_t)p->fft_size, p->bins[c], bins);
        if (p->tr_spec) fvec_push(&p->spec_trace[c], p->bins[c], bins);
    }
    for (int k = 0; k < p->n_machines; k++) {
        vadm *m = &p->machines[k];
        float min_volume = 999, max_volume = 0;                    /* VADMachine.zig:134-146 */
        for (int c = 0; c < p->n_channels; c++) {
            float v = 0.0f;                                        /* PipelineFFT.zig:99-112: a SUM */
            for (long i = m->min_bin; i < m->max_bin + 1; i++) v += p->bins[c][i];
            if (v < min_volume) min_volume = v;
            if (v > max_volume) max_volume = v;
        }
        uint64_t sn;
        int rec = vadm_step(m, min_volume, volume_ratio, has_vad ? vad : 0.0f, index, &sn);
        if (k == 0) {
            if (rec) {
                if (p->n_events == p->cap_events) { p->cap_events = p->cap_events ? 2 * p->cap_events : 64; p->events = (orc_capture_event *)realloc(p->events, p->cap_events * sizeof(*p->events)); }
                orc_capture_event ev = { index, rec, sn };
                p->events[p->n_events++] = ev;
            }
            if (p->tr_rec) {
                if (p->n_recs == p->cap_recs) { p->cap_recs = p->cap_recs ? 2 * p->cap_recs : 1024; p->recs = (orc_stft_record *)realloc(p->recs, p->cap_recs * sizeof(*p->recs)); }
                orc_stft_record r = { index, min_volume, volume_ratio, has_vad ? vad : 0.0f, has_vad ? 1.0f : 0.0f };
                p->recs[p->n_recs++] = r;
            }
        }
    }
}

/* preAnalyzeSegment (VAD.zig:253-272) on n samples per channel */
static float pre_analyze(orc_pipeline *p, float *const *ch, size_t n) {
    float vol_min = 1, vol_max = 0;
    for (int c = 0; c < p->n_channels; c++) {
        float vol = orc_rms_volume(ch[c], n, NULL, 0);
        if (vol < vol_min) vol_min = vol;
        if (vol > vol_max) vol_max = vol;
    }
    return vol_max == 0 ? 0 : vol_min / vol_max;
}

/* collectInputStep (VAD.zig:214-251) */
static void run_vad(orc_pipeline *p) {
    size_t frame = p->use_denoiser ? 480 : (size_t)p->fft_size;
    int C = p->n_channels;
    float *in[64];
    while (p->total_write_count - p->read_count >= frame) {
        uint64_t from = p->read_count;
        p->read_count = from + frame;
        for (int c = 0; c < C; c++) in[c] = p->pending[c].v + p->pending_off;
        p->pending_off += frame;
        float ratio = pre_analyze(p, in, frame);
        if (!p->use_denoiser) { fft_step(p, in, from, 0, 0, ratio); continue; }
        /* denoiserStep (VAD.zig:274-296) */
        float vad_low = 1;
        for (int c = 0; c < C; c++) {
            float vad = orc_denoiser_denoise(p->denoiser, in[c], p->denoised[c]);
            if (vad < vad_low) vad_low = vad;
            if (p->tr_pcm) fvec_push(&p->pcm_trace[c], p->denoised[c], 480);
        }
        if (p->tr_vad) fvec_push(&p->vad_trace, &vad_low, 1);
        /* fftBufferStep (VAD.zig:298-348) + SegmentWriter.write (SegmentWriter.zig:40-108) */
        size_t off = 0, buf_len = (size_t)p->fft_size;
        for (;;) {
            size_t remaining = p->fft_write < buf_len ? buf_len - p->fft_write : 0;
            size_t written = remaining < 480 - off ? remaining : 480 - off;
            for (int c = 0; c < C; c++) memcpy(p->fft_buf[c] + p->fft_write, p->denoised[c] + off, written * sizeof(float));
            p->fft_write += written;
            off += written;
            int full = p->fft_write == buf_len;
            float share = (float)written / (float)buf_len;
            p->tmp_rnn_vad += vad_low * share;
            p->tmp_vol_ratio += ratio * share;
            if (full) {
                fft_step(p, p->fft_buf, p->fft_index, 1, vad_low, p->tmp_vol_ratio);
                p->fft_index += buf_len;
                p->fft_write = 0;
                p->tmp_rnn_vad = 0; p->tmp_vol_ratio = 0;
            }
            if (off == 480) break;
        }
    }
    /* drop consumed samples */
    if (p->pending_off > (1u << 20) || p->pending_off == p->pending[0].n) {
        for (int c = 0; c < C; c++) {
            memmove(p->pending[c].v, p->pending[c].v + p->pending_off, (p->pending[c].n - p->pending_off) * sizeof(float));
            p->pending[c].n -= p->pending_off;
        }
        p->pending_off = 0;
    }
}

uint64_t orc_pipeline_push(orc_pipeline *p, const float *const *channel_pcm, size_t n) {   /* AudioPipeline.zig:86-120 */
    uint64_t first = p->total_write_count;
    /* the reference writes in chunks of capacity/2 and runs the VAD after each; frame
     * boundaries depend only on the running sample count, so chunking is invisible */
    for (int c = 0; c < p->n_channels; c++) fvec_push(&p->pending[c], channel_pcm[c], n);
    p->total_write_count += n;
    run_vad(p);
    return first;
}

size_t orc_pipeline_segments(const orc_pipeline *p, int k, orc_segment *out, size_t cap) {
    if (k < 0 || k >= p->n_machines) return 0;
    size_t n = p->machines[k].n_segs;
    for (size_t i = 0; i < n && i < cap; i++) out[i] = p->machines[k].segs[i];
    return n;
}
size_t orc_pipeline_events(const orc_pipeline *p, orc_capture_event *out, size_t cap) {
    for (size_t i = 0; i < p->n_events && i < cap; i++) out[i] = p->events[i];
    return p->n_events;
}
size_t orc_pipeline_stft_records(const orc_pipeline *p, orc_stft_record *out, size_t cap) {
    for (size_t i = 0; i < p->n_recs && i < cap; i++) out[i] = p->recs[i];
    return p->n_recs;
}
size_t orc_pipeline_denoised(const orc_pipeline *p, int c, float *out, size_t cap) {
    size_t n = p->pcm_trace[c].n;
    memcpy(out, p->pcm_trace[c].v, (n < cap ? n : cap) * sizeof(float));
    return n;
}
size_t orc_pipeline_rnn_vad(const orc_pipeline *p, float *out, size_t cap) {
    size_t n = p->vad_trace.n;
    memcpy(out, p->vad_trace.v, (n < cap ? n : cap) * sizeof(float));
    return n;
}
size_t orc_pipeline_spectra(const orc_pipeline *p, int c, float *out, size_t cap) {
    size_t n = p->spec_trace[c].n;
    memcpy(out, p->spec_trace[c].v, (n < cap ? n : cap) * sizeof(float));
    return n;
}
const char *orc_build_kind(void) { return leaf_kind(); }
