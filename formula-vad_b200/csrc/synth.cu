// synth.cu -- seeded synthetic 48 kHz "onboard radio" input, the same bits on
// host and device: integer hashing/phase arithmetic, one 4096-entry sine table computed once on the
// host, and f32 operations that are individually rounded on both sides (-fmad=false / -ffp-contract=off).
// Every sample is a pure function of (seed, stream, n), so any range can be generated in parallel.
//
// Content per stream: an "engine" bed (8 harmonics of a slowly wandering 80-300 Hz fundamental plus
// white noise, about -20 dBFS), "radio speech" bursts (harmonic stack, f0 100-250 Hz, syllabic 3-5 Hz
// envelope, 0.3-8 s long, at most one per 16 s slot), exact-zero gaps (digital silence: RNNoise's
// silence gate denoise.c:358-362 and vol_max==0 in VAD.zig:267) and very quiet gaps.
#include "common.cuh"
#include "../../include/fvad_synth_core.h"
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

namespace fvad {

constexpr int kSinN = FVAD_SYNTH_SIN_N;

static std::vector<float> g_sin_host;
static float *g_sin_dev[64];
static std::mutex g_mu;

static const float *sin_host() {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_sin_host.empty()) {
        g_sin_host.resize(kSinN);
        for (int i = 0; i < kSinN; i++) g_sin_host[i] = (float)sin(FVAD_SYNTH_TWO_PI * i / kSinN);
    }
    return g_sin_host.data();
}

__global__ void k_synth(float *dst, size_t stream_stride, const float *tab, uint64_t seed, uint64_t first_stream,
                        size_t n_streams, uint64_t first_sample, size_t n) {
    const size_t s = blockIdx.y;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[s * stream_stride + i] = fvad_synth_sample(tab, seed, first_stream + s, first_sample + i);
}

}  // namespace fvad

using namespace fvad;

extern "C" int fvad_synth_fill_host(float *dst, uint64_t seed, uint64_t stream_id, uint64_t first_sample, size_t n) {
    if (!dst && n) return set_error(FVAD_EINVALID_ARGUMENT, "null dst");
    const float *tab = sin_host();
    for (size_t i = 0; i < n; i++) dst[i] = fvad_synth_sample(tab, seed, stream_id, first_sample + i);
    return FVAD_OK;
}

extern "C" int fvad_synth_fill_device(float *d_dst, size_t stream_stride, uint64_t seed, uint64_t first_stream_id,
                                      size_t n_streams, uint64_t first_sample, size_t n, void *cuda_stream) {
    if (!d_dst && n) return set_error(FVAD_EINVALID_ARGUMENT, "null dst");
    if (!n || !n_streams) return FVAD_OK;
    int device = 0;
    FVAD_CUDA_TRY(cudaGetDevice(&device));
    const float *tab = sin_host();
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_sin_dev[device]) {
            FVAD_CUDA_TRY(cudaMalloc(&g_sin_dev[device], sizeof(float) * kSinN));
            FVAD_CUDA_TRY(cudaMemcpy(g_sin_dev[device], tab, sizeof(float) * kSinN, cudaMemcpyHostToDevice));
        }
    }
    for (size_t s0 = 0; s0 < n_streams; s0 += 65535) {
        const size_t ns = std::min<size_t>(65535, n_streams - s0);
        unsigned gx = (unsigned)std::min<size_t>(ceil_div(n, 256 * 4), 2048);
        FVAD_LAUNCH(k_synth, dim3(gx, (unsigned)ns), 256, 0, (cudaStream_t)cuda_stream, d_dst + s0 * stream_stride,
                    stream_stride, g_sin_dev[device], seed, first_stream_id + s0, ns, first_sample, n);
    }
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}
