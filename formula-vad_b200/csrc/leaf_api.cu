// leaf_api.cu -- C ABI of the leaf drop-ins: FFT.zig (single transform and batched STFT) and the
// VADMachine sweep over shared features.
#include "common.cuh"
#include "stft.cuh"
#include "vadm.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

using namespace fvad;

namespace {
struct SweepBuffers {
    void *p[8] = {};
    int n = 0;
    template <class T> int get(T **out, size_t count) {
        *out = nullptr;
        cudaError_t e = cudaMalloc((void **)out, sizeof(T) * (count ? count : 1));
        if (e != cudaSuccess)
            return set_error(e == cudaErrorMemoryAllocation ? FVAD_EOUT_OF_MEMORY : FVAD_ECUDA, "cudaMalloc(%zu bytes): %s",
                             sizeof(T) * count, cudaGetErrorString(e));
        p[n++] = *out;
        return FVAD_OK;
    }
    ~SweepBuffers() { for (int i = 0; i < n; i++) cudaFree(p[i]); }
};
struct LeafDeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit LeafDeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~LeafDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
}  // namespace

struct fvad_fft {
    size_t n_fft, sample_rate;
    int device;
    const StftTables *tabs;
    float *d_in = nullptr, *d_window = nullptr, *d_out = nullptr, *d_band = nullptr;
    std::vector<float> h_in, h_window;
};

extern "C" int fvad_fft_create(size_t n_fft, size_t sample_rate, fvad_fft **out) {
    if (!out) return set_error(FVAD_EINVALID_ARGUMENT, "null out");
    *out = nullptr;
    if (n_fft == 0 || (n_fft % 2) != 0) return set_error(FVAD_EINVALID_FFT_SIZE, "n_fft %zu must be even and non-zero", n_fft);   // FFT.zig:29-31
    if (n_fft != (size_t)kFft) FVAD_TRY(stft_generic_supported(n_fft));     // KissFFT's generic plan (stft_generic.cu)
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    fvad_fft *f = new (std::nothrow) fvad_fft();
    if (!f) return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed");
    f->n_fft = n_fft; f->sample_rate = sample_rate; f->device = device;
    int rc = n_fft == (size_t)kFft ? stft_tables_get(device, &f->tabs) : FVAD_OK;
    if (rc != FVAD_OK) { delete f; return rc; }
    if (cudaMalloc(&f->d_in, sizeof(float) * n_fft) != cudaSuccess || cudaMalloc(&f->d_window, sizeof(float) * n_fft) != cudaSuccess ||
        cudaMalloc(&f->d_out, sizeof(float) * (n_fft / 2 + 1)) != cudaSuccess || cudaMalloc(&f->d_band, sizeof(float) * kMaxBands) != cudaSuccess) {
        fvad_fft_destroy(f);
        return set_error(FVAD_EOUT_OF_MEMORY, "device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    f->h_in.resize(n_fft); f->h_window.resize(n_fft);
    *out = f;
    return FVAD_OK;
}

extern "C" void fvad_fft_destroy(fvad_fft *f) {
    if (!f) return;
    cudaFree(f->d_in); cudaFree(f->d_window); cudaFree(f->d_out); cudaFree(f->d_band);
    delete f;
}

extern "C" size_t fvad_fft_bin_count(const fvad_fft *f) { return f->n_fft / 2 + 1; }                       // FFT.zig:101-103
extern "C" float fvad_fft_bin_width(const fvad_fft *f) { return (float)f->sample_rate / (float)f->n_fft; } // FFT.zig:106-111
extern "C" float fvad_fft_nyquist_freq(const fvad_fft *f) { return (float)f->sample_rate / 2; }            // FFT.zig:114-117
extern "C" int fvad_fft_freq_to_bin(const fvad_fft *f, float freq, size_t *bin) {                          // FFT.zig:120-131
    if (freq > fvad_fft_nyquist_freq(f)) return set_error(FVAD_EOUT_OF_RANGE, "frequency above Nyquist");
    if (freq < 0) return set_error(FVAD_ENEGATIVE_FREQUENCY, "negative frequency");
    *bin = (size_t)roundf(freq / fvad_fft_bin_width(f));
    return FVAD_OK;
}
extern "C" int fvad_fft_bin_to_freq(const fvad_fft *f, size_t bin, float *freq) {                          // FFT.zig:134-144
    if (bin > fvad_fft_bin_count(f) - 1) return set_error(FVAD_EOUT_OF_RANGE, "bin out of range");
    *freq = (float)bin * fvad_fft_bin_width(f);
    return FVAD_OK;
}
extern "C" void fvad_hann_window_periodic(float *w, size_t n) { host_hann_window_periodic(w, n); }
extern "C" float fvad_window_norm_factor(const float *w, size_t n) { return host_window_norm_factor(w, n); }

// FFT.fft (src/FFT.zig:70-98): host SplitSlice in, normalised magnitudes out.
extern "C" int fvad_fft_run(fvad_fft *f, const float *first, size_t n_first, const float *second, size_t n_second,
                            const float *window, size_t n_window, float *result, size_t n_result) {
    if (!f) return set_error(FVAD_EINVALID_ARGUMENT, "null fft");
    if (n_first + n_second != f->n_fft) return set_error(FVAD_EINVALID_SAMPLES_LEN, "samples length %zu != n_fft", n_first + n_second);
    if (n_window != f->n_fft) return set_error(FVAD_EINVALID_WINDOW_LEN, "window length %zu != n_fft", n_window);
    if (n_result != f->n_fft / 2 + 1) return set_error(FVAD_EINVALID_RESULT_LEN, "result length %zu != bin count", n_result);
    if (n_first) memcpy(f->h_in.data(), first, sizeof(float) * n_first);
    if (n_second) memcpy(f->h_in.data() + n_first, second, sizeof(float) * n_second);
    // window norm factor is recomputed per call as the reference does (FFT.zig:94); in-order f32 sum
    const float norm = host_window_norm_factor(window, n_window) / (float)(f->n_fft / 2);
    LeafDeviceGuard g(f->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select CUDA device %d (no CPU fallback exists)", f->device);
    FVAD_CUDA_TRY(cudaMemcpyAsync(f->d_in, f->h_in.data(), sizeof(float) * f->n_fft, cudaMemcpyHostToDevice, 0));
    FVAD_CUDA_TRY(cudaMemcpyAsync(f->d_window, window, sizeof(float) * f->n_fft, cudaMemcpyHostToDevice, 0));
    StftBands b{}; b.n_bands = 0;
    if (f->n_fft == (size_t)kFft)
        FVAD_TRY(stft_launch(*f->tabs, f->d_in, 1, f->n_fft, 1, f->n_fft, b, f->d_band, 1, 0, f->d_out, f->d_window, norm, f->device, 0));
    else
        FVAD_TRY(stft_generic_launch(f->device, f->n_fft, f->d_in, 1, f->n_fft, 1, f->n_fft, b, f->d_band, 1, 0, f->d_out, f->d_window, norm, 0));
    FVAD_CUDA_TRY(cudaMemcpy(result, f->d_out, sizeof(float) * n_result, cudaMemcpyDeviceToHost));
    return FVAD_OK;
}

extern "C" int fvad_stft_batch(fvad_fft *f, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames,
                               size_t hop, size_t min_bin, size_t max_bin, float *d_band_sum, float *d_spectra,
                               void *cuda_stream) {
    if (!f || !d_pcm || !d_band_sum) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_stft_batch: null argument");
    if (max_bin >= f->n_fft / 2 + 1 || min_bin > max_bin) return set_error(FVAD_EOUT_OF_RANGE, "band %zu..%zu out of range", min_bin, max_bin);
    StftBands b{}; b.n_bands = 1; b.lo[0] = (int)min_bin; b.hi[0] = (int)max_bin;
    if (f->n_fft != (size_t)kFft)
        return stft_generic_launch(f->device, f->n_fft, d_pcm, n_rows, row_stride, n_frames, hop, b, d_band_sum, n_frames, 0, d_spectra,
                                   nullptr, 0.f, (cudaStream_t)cuda_stream);
    return stft_launch(*f->tabs, d_pcm, n_rows, row_stride, n_frames, hop, b, d_band_sum, n_frames, 0, d_spectra, nullptr, 0.f,
                       f->device, (cudaStream_t)cuda_stream);
}

// ---- Spectrogram.compute (src/audio_utils/Spectrogram.zig:30-94) ---------------------------------------------
extern "C" int fvad_spectrogram_shape(size_t n_samples, size_t sample_rate, size_t fft_size, size_t hop_size,
                                      fvad_spectrogram_info *info) {
    if (!info) return set_error(FVAD_EINVALID_ARGUMENT, "null info");
    if (fft_size == 0 || (fft_size % 2) != 0) return set_error(FVAD_EINVALID_FFT_SIZE, "fft_size %zu must be even and non-zero", fft_size);
    if (hop_size == 0) return set_error(FVAD_EINVALID_ARGUMENT, "hop_size 0 (division by zero in the reference)");
    if (n_samples < fft_size) return set_error(FVAD_EINSUFFICIENT_SAMPLES, "%zu samples < fft_size %zu", n_samples, fft_size);   // :34-36
    if (hop_size > n_samples) return set_error(FVAD_EOUT_OF_RANGE, "hop_size %zu > %zu samples (usize underflow in the reference)", hop_size, n_samples);
    const size_t steps = (n_samples - hop_size) / hop_size;                                                                          // :39
    if (steps > 0 && (steps - 1) * hop_size + fft_size > n_samples)
        return set_error(FVAD_EOUT_OF_RANGE, "last frame ends at %zu > %zu samples (the reference's slice panics)",
                         (steps - 1) * hop_size + fft_size, n_samples);
    info->height = fft_size / 2 + 1;
    info->width = steps;
    info->bin_width = (float)sample_rate / (float)fft_size;
    info->nyquist_freq = (float)sample_rate / 2;
    const size_t processed = steps ? (steps - 1) * hop_size + fft_size : 0;                                                          // :66
    info->length_sec = (double)processed / (double)sample_rate;
    return FVAD_OK;
}

extern "C" int fvad_spectrogram_compute(const float *samples, size_t n_samples, size_t sample_rate, size_t fft_size,
                                        size_t hop_size, float *values, size_t values_len, float *bin_labels,
                                        fvad_spectrogram_info *info) {
    fvad_spectrogram_info local;
    if (!info) info = &local;
    FVAD_TRY(fvad_spectrogram_shape(n_samples, sample_rate, fft_size, hop_size, info));
    if (!samples || !values) return set_error(FVAD_EINVALID_ARGUMENT, "null samples/values");
    if (values_len != info->width * info->height)
        return set_error(FVAD_EINVALID_RESULT_LEN, "values length %zu != width*height %zu", values_len, info->width * info->height);
    if (fft_size != (size_t)kFft) FVAD_TRY(stft_generic_supported(fft_size));
    if (fft_size == (size_t)kFft && (hop_size & 1)) return set_error(FVAD_EUNSUPPORTED, "odd hop_size %zu: the 2048-point kernel loads sample pairs", hop_size);
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    const StftTables *tabs = nullptr;
    if (fft_size == (size_t)kFft) FVAD_TRY(stft_tables_get(device, &tabs));
    if (bin_labels)
        for (size_t k = 0; k < info->height; k++) bin_labels[k] = (float)k * info->bin_width;     // FFT.binToFreq, FFT.zig:134-144
    if (info->width == 0) return FVAD_OK;
    float *d_in = nullptr, *d_out = nullptr;
    int rc = FVAD_OK;
    if (cudaMalloc(&d_in, sizeof(float) * n_samples) != cudaSuccess || cudaMalloc(&d_out, sizeof(float) * values_len) != cudaSuccess)
        rc = set_error(FVAD_EOUT_OF_MEMORY, "device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc == FVAD_OK && cudaMemcpy(d_in, samples, sizeof(float) * n_samples, cudaMemcpyHostToDevice) != cudaSuccess)
        rc = set_error(FVAD_ECUDA, "H2D: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc == FVAD_OK) {
        StftBands b{}; b.n_bands = 0;
        // one row, `width` overlapping frames `hop_size` apart; the window norm factor FFT.fft recomputes per call
        // (FFT.zig:94) is the table's (same in-order f32 sum over the same Hann values)
        rc = fft_size == (size_t)kFft
                 ? stft_launch(*tabs, d_in, 1, 0 /* one row: stride unused */, info->width, hop_size, b, nullptr, info->width, 0, d_out, nullptr, 0.f, device, 0)
                 : stft_generic_launch(device, fft_size, d_in, 1, 0, info->width, hop_size, b, nullptr, info->width, 0, d_out, nullptr, 0.f, 0);
    }
    if (rc == FVAD_OK && cudaMemcpy(values, d_out, sizeof(float) * values_len, cudaMemcpyDeviceToHost) != cudaSuccess)
        rc = set_error(FVAD_ECUDA, "D2H: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(d_in); cudaFree(d_out);
    return rc;
}

// ---- VADMachine sweep (BASELINE config 5) --------------------------------------------------
// One scan per (stream, config) over the same records: VAD.stateMachineStep's alt machines
// (VAD.zig:375-380) run on "shared features".  Host records in, host lists out; the scans are vadm_sweep.cu's.
// A record's vad is `fft_input.vad orelse 0` (VADMachine.zig:240,245): has_vad == 0 contributes 0.  The index of record f
// is taken as records[s][0].index + f * fft_size; records that say otherwise are rejected.

extern "C" int fvad_vadm_sweep(const fvad_stft_record *records, size_t n_streams, size_t n_records,
                               const fvad_vad_machine_config *configs, size_t n_configs, size_t sample_rate,
                               size_t fft_size, int device, fvad_segment *segs, size_t max_segs, uint32_t *counts) {
    if (!records || !configs || !segs || !counts || !n_streams || !n_configs || !max_segs)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_vadm_sweep: null/zero argument");
    if (device < 0) {
        cudaError_t e = cudaGetDevice(&device);
        if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    }
    LeafDeviceGuard g(device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select CUDA device %d (no CPU fallback exists)", device);
    std::vector<VadmParams> params(n_configs);
    for (size_t k = 0; k < n_configs; k++) {
        size_t nd = 0;
        FVAD_TRY(vadm_params_from_config(configs[k], sample_rate, fft_size, &params[k], &nd));
        params[k].band = 0;                                   // the records carry ONE min_volume: the band they were made with
    }
    int max_long = 0, max_short = 0, max_ratio = 0;
    vadm_sweep_rows(params.data(), n_configs, &max_long, &max_short, &max_ratio);
    const size_t n_scans = n_streams * n_configs, total = n_streams * n_records;
    memset(counts, 0, sizeof(uint32_t) * n_scans);
    if (!n_records) return FVAD_OK;
    // split the records into the planes the scans read
    std::vector<float> minv(total), ratio(total), vad(total);
    for (size_t s = 0; s < n_streams; s++) {
        const fvad_stft_record *r = records + s * n_records;
        if (r[0].index != records[0].index)
            return set_error(FVAD_EINVALID_ARGUMENT, "fvad_vadm_sweep: stream %zu starts at index %llu, stream 0 at %llu", s,
                             (unsigned long long)r[0].index, (unsigned long long)records[0].index);
        for (size_t f = 0; f < n_records; f++) {
            if (r[f].index != r[0].index + f * fft_size)
                return set_error(FVAD_EINVALID_ARGUMENT, "fvad_vadm_sweep: record %zu of stream %zu has index %llu, expected %llu "
                                 "(records must be consecutive STFT frames)", f, s, (unsigned long long)r[f].index,
                                 (unsigned long long)(r[0].index + f * fft_size));
            minv[s * n_records + f] = r[f].min_volume; ratio[s * n_records + f] = r[f].vol_ratio;
            vad[s * n_records + f] = r[f].has_vad != 0.0f ? r[f].rnn_vad : 0.0f;
        }
    }
    SweepBuffers buf;
    VadmParams *d_params; double *d_rings; float *d_minv, *d_ratio, *d_vad; fvad_segment *d_segs; uint32_t *d_counts;
    FVAD_TRY(buf.get(&d_params, n_configs));
    FVAD_TRY(buf.get(&d_rings, (size_t)(max_long + max_short + max_ratio) * n_scans));
    FVAD_TRY(buf.get(&d_minv, total)); FVAD_TRY(buf.get(&d_ratio, total)); FVAD_TRY(buf.get(&d_vad, total));
    FVAD_TRY(buf.get(&d_segs, n_scans * max_segs)); FVAD_TRY(buf.get(&d_counts, n_scans));
    FVAD_CUDA_TRY(cudaMemcpy(d_params, params.data(), sizeof(VadmParams) * n_configs, cudaMemcpyHostToDevice));
    FVAD_CUDA_TRY(cudaMemcpy(d_minv, minv.data(), sizeof(float) * total, cudaMemcpyHostToDevice));
    FVAD_CUDA_TRY(cudaMemcpy(d_ratio, ratio.data(), sizeof(float) * total, cudaMemcpyHostToDevice));
    FVAD_CUDA_TRY(cudaMemcpy(d_vad, vad.data(), sizeof(float) * total, cudaMemcpyHostToDevice));
    VadmSweepIo io{};
    io.min_volume = d_minv; io.band_stride = 0; io.vol_ratio = d_ratio; io.vad = d_vad;
    io.rec_stride = n_records; io.n_records = n_records;
    io.first_index = records[0].index; io.index_step = fft_size;
    io.segs = d_segs; io.max_segs = max_segs; io.counts = d_counts; io.fallbacks = nullptr;
    FVAD_TRY(vadm_sweep_launch(d_params, d_rings, max_long, max_short, n_streams, (int)n_configs, io, 0));
    FVAD_CUDA_TRY(cudaMemcpy(counts, d_counts, sizeof(uint32_t) * n_scans, cudaMemcpyDeviceToHost));
    FVAD_CUDA_TRY(cudaMemcpy(segs, d_segs, sizeof(fvad_segment) * n_scans * max_segs, cudaMemcpyDeviceToHost));
    return FVAD_OK;
}
