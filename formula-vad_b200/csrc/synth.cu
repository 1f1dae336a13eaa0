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
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

namespace fvad {

constexpr int kSinBits = 12, kSinN = 1 << kSinBits;
constexpr uint64_t kSlot = 16ull * 48000ull;          // one scheduling slot = 16 s

__host__ __device__ inline uint64_t mix64(uint64_t z) {     // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ inline uint64_t hash3(uint64_t seed, uint64_t a, uint64_t b) {
    return mix64(mix64(seed ^ (a * 0xD6E8FEB86659FD93ull)) ^ (b * 0xA0761D6478BD642Full));
}
// sine of a 32-bit phase (one turn = 2^32) by linear interpolation in the table
__host__ __device__ inline float sin_u32(const float *tab, uint32_t phase) {
    const uint32_t idx = phase >> (32 - kSinBits);
    const float frac = (float)(phase & ((1u << (32 - kSinBits)) - 1)) * (1.0f / (float)(1u << (32 - kSinBits)));
    const float s0 = tab[idx], s1 = tab[(idx + 1) & (kSinN - 1)];
    return s0 + (s1 - s0) * frac;
}
// phase increment per sample for a frequency in milli-hertz
__host__ __device__ inline uint32_t inc_mhz(uint64_t mhz) { return (uint32_t)((mhz << 32) / 48000000ull); }

__host__ __device__ inline float synth_sample(const float *tab, uint64_t seed, uint64_t stream, uint64_t n) {
    const uint64_t hs = hash3(seed, stream, 0x5EEDull);
    // ---- gaps -------------------------------------------------------------------------------
    if ((stream & 1) && n < 24000) return 0.0f;                         // half a second of digital silence
    const uint64_t slot = n / kSlot, in_slot = n - slot * kSlot;
    const uint64_t hk = hash3(seed, stream, 0x1000 + slot);
    const uint32_t gap_kind = (uint32_t)(hk & 7);                       // 0: exact zeros, 1: very quiet
    const uint64_t gap_len = 48000ull + (uint64_t)((hk >> 3) % 96000ull);   // 1..3 s at the slot start
    float level = 1.0f;
    if (in_slot < gap_len) {
        if (gap_kind == 0) return 0.0f;
        if (gap_kind == 1) level = 1.0e-4f;
    }
    // ---- engine bed ---------------------------------------------------------------------------
    const uint64_t f0_mhz = 80000ull + (hs % 220000ull);
    const uint32_t lfo = inc_mhz(50 + ((hs >> 20) % 200)) * (uint32_t)n;      // 0.05-0.25 Hz wander
    const float wander = sin_u32(tab, lfo);
    // phase modulation: up to +-12 turns of the fundamental -> a few Hz to tens of Hz of deviation
    const int64_t pm = (int64_t)(wander * (12.0f * 4294967296.0f));
    const uint32_t ph = (uint32_t)((uint64_t)inc_mhz(f0_mhz) * n + (uint64_t)pm);
    float v = 0.0f;
#pragma unroll
    for (int h = 1; h <= 8; h++) {
        const float a = 0.003f / (float)h;
        v = v + a * sin_u32(tab, ph * (uint32_t)h + (uint32_t)(hs >> (4 * h)) * 0x9E3779B9u);
    }
    const uint64_t hn = hash3(seed, stream ^ 0xABCDEF, n);
    v = v + (float)(int32_t)(uint32_t)hn * (0.03f / 2147483648.0f);    // white noise, ~-35 dBFS rms
    // ---- radio speech burst ---------------------------------------------------------------------
    if (((hk >> 24) & 3) != 0) {                                         // 3 slots out of 4 carry a burst
        const uint64_t dur = 14400ull + (uint64_t)((hk >> 26) % 369600ull);       // 0.3 .. 8 s
        const uint64_t start = 4ull * 48000ull + (uint64_t)((hk >> 45) % (kSlot - 4ull * 48000ull - dur));
        if (in_slot >= start && in_slot < start + dur) {
            const uint64_t t = in_slot - start;
            const uint64_t sf0 = 100000ull + ((hk >> 8) % 150000ull);
            const uint32_t sph = inc_mhz(sf0) * (uint32_t)t;
            const uint32_t syl = inc_mhz(3000 + ((hk >> 12) % 2000)) * (uint32_t)t;
            float env = 0.5f - 0.5f * sin_u32(tab, syl + 0x40000000u);   // 0..1, starts at 0
            env = 0.15f + 0.85f * env;
            // 10 ms linear ramps at both ends
            const uint64_t edge = t < dur - t ? t : dur - t;
            if (edge < 480) env = env * ((float)edge * (1.0f / 480.0f));
            float sp = 0.0f;
#pragma unroll
            for (int h = 1; h <= 10; h++) {
                const uint32_t wsel = (uint32_t)(hk >> (2 * h)) & 3u;
                const float wgt = (wsel == 0 ? 0.25f : wsel == 1 ? 0.5f : wsel == 2 ? 0.75f : 1.0f) * (h <= 6 ? 0.12f : 0.05f);
                sp = sp + wgt * sin_u32(tab, sph * (uint32_t)h);
            }
            v = v + env * sp;
        }
    }
    v = v * level;
    if (v > 1.0f) v = 1.0f;
    if (v < -1.0f) v = -1.0f;
    return v;
}

static std::vector<float> g_sin_host;
static float *g_sin_dev[64];
static std::mutex g_mu;

static const float *sin_host() {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_sin_host.empty()) {
        g_sin_host.resize(kSinN);
        for (int i = 0; i < kSinN; i++) g_sin_host[i] = (float)sin(2.0 * 3.14159265358979323846 * i / kSinN);
    }
    return g_sin_host.data();
}

__global__ void k_synth(float *dst, size_t stream_stride, const float *tab, uint64_t seed, uint64_t first_stream,
                        size_t n_streams, uint64_t first_sample, size_t n) {
    const size_t s = blockIdx.y;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[s * stream_stride + i] = synth_sample(tab, seed, first_stream + s, first_sample + i);
}

}  // namespace fvad

using namespace fvad;

extern "C" int fvad_synth_fill_host(float *dst, uint64_t seed, uint64_t stream_id, uint64_t first_sample, size_t n) {
    if (!dst && n) return set_error(FVAD_EINVALID_ARGUMENT, "null dst");
    const float *tab = sin_host();
    for (size_t i = 0; i < n; i++) dst[i] = synth_sample(tab, seed, stream_id, first_sample + i);
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
