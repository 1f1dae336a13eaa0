// stft.cuh -- batched windowed real STFT (2048 points) + speech-band reduction.
// Replaces FFT.fft / PipelineFFT.fft / averageVolumeInBand (src/FFT.zig:70-98,147-177;
// src/AudioPipeline/PipelineFFT.zig:88-112) over lib/kissfft (kiss_fftr.c:63-117,
// kiss_fft.c:38-84,235-300) with one kernel over (rows x frames).
#pragma once
#include "common.cuh"

namespace fvad {

constexpr int kFft = 2048;          // the size this build's kernel covers (VAD.Config default, VAD.zig:18)
constexpr int kFftC = kFft / 2;     // complex length handed to kiss_fft
constexpr int kBins = kFft / 2 + 1;
constexpr int kMaxBands = 8;

struct StftBands {
    int n_bands;
    int lo[kMaxBands], hi[kMaxBands];   // inclusive bin ranges (freqToBin(min), freqToBin(max))
};

struct StftTwA {
    float2 v[9];                        // stage m=4 twiddles tw[k*64*q], k,q = 1..3 (uniform over threads)
    unsigned long long neg_zero;        // the pair (-0.0f, -0.0f): addend of the product-forming fma (see prod2)
};

struct StftTables {                  // per device
    float *d_window = nullptr;       // periodic Hann, 2048 (PipelineFFT.zig:66-68)
    float2 *d_twB4 = nullptr;        // [15][64] per-thread twiddles of stages m=16 and m=64
    float2 *d_tw4 = nullptr;         // 1024 twiddles (kiss_fft.c:357-363)
    float2 *d_stw4 = nullptr;        // 512 super twiddles (kiss_fftr.c:53-59)
    StftTwA twA{};
    float norm_factor = 0;           // (N / sum(w)) / (N/2)  (FFT.zig:94,165)
};

// Host-side tables, computed with the same libm the reference's C uses.
void host_hann_window_periodic(float *w, size_t n);
float host_window_norm_factor(const float *w, size_t n);
int stft_tables_get(int device, const StftTables **out);

// frame f of row r = d_pcm[r*row_stride + f*hop ...+2048).  Outputs:
//   d_band[(band*n_rows + r)*band_frame_stride + band_frame_offset + f]  sum of bins lo..hi
//   d_spectra (optional) [(r*n_frames + f)*1025]
// If d_window_override != nullptr it is used instead of the Hann table (FFT.fft takes the
// window as an argument) with the given norm factor.
int stft_launch(const StftTables &t, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames,
                size_t hop, const StftBands &bands, float *d_band, size_t band_frame_stride,
                size_t band_frame_offset, float *d_spectra, const float *d_window_override,
                float norm_override, int device, cudaStream_t stream);

// Any other even n_fft whose complex length n_fft/2 is made of the factors 2, 3, 4, 5 (stft_generic.cu): KissFFT's plan
// executed stage by stage, one CTA per frame.  Same argument meaning as stft_launch; the Hann window / norm factor of the
// size are built on first use.
int stft_generic_supported(size_t n_fft);                 // FVAD_OK, FVAD_EINVALID_FFT_SIZE or FVAD_EUNSUPPORTED (message set)
int stft_generic_norm_factor(int device, size_t n_fft, float *norm);
int stft_generic_launch(int device, size_t n_fft, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames, size_t hop,
                        const StftBands &bands, float *d_band, size_t band_frame_stride, size_t band_frame_offset, float *d_spectra,
                        const float *d_window_override, float norm_override, cudaStream_t stream);

}  // namespace fvad
