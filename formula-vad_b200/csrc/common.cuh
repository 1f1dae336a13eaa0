// common.cuh -- shared host/device plumbing for libfvad_cuda.so (sm_100a only).
// The whole library is compiled with -fmad=false: every f32/f64 product and sum is a
// separately rounded IEEE operation, which is what makes the results bit-comparable
// with the reference's C arithmetic (DESIGN.md "Exactness").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/fvad.h"

namespace fvad {

int set_error(int code, const char *fmt, ...);
extern std::atomic<uint64_t> g_launch_count;

#define FVAD_CUDA_TRY(expr)                                                                         \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            return ::fvad::set_error(FVAD_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr,      \
                                     cudaGetErrorString(e__));                                      \
    } while (0)

#define FVAD_TRY(expr)                    \
    do {                                  \
        int rc__ = (expr);                \
        if (rc__ != FVAD_OK) return rc__; \
    } while (0)

// Optional per-kernel timing (fvad_profile_enable): CUDA events recorded on the launching stream
// around every launch, aggregated by kernel name in fvad_profile_report.
extern std::atomic<int> g_profile_on;
void profile_begin(const char *name, cudaStream_t stream);
void profile_end(cudaStream_t stream);

// Every kernel launch of the library goes through this so that fvad_launch_count() is honest.
#define FVAD_LAUNCH(kernel, grid, block, smem, stream, ...) \
    FVAD_LAUNCH_NAMED(#kernel, kernel, grid, block, smem, stream, __VA_ARGS__)
#define FVAD_LAUNCH_NAMED(name, kernel, grid, block, smem, stream, ...)     \
    do {                                                                    \
        const bool prof__ = ::fvad::g_profile_on.load(std::memory_order_relaxed) != 0; \
        if (prof__) ::fvad::profile_begin(name, (stream));                  \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);         \
        if (prof__) ::fvad::profile_end((stream));                          \
        ::fvad::g_launch_count.fetch_add(1, std::memory_order_relaxed);     \
    } while (0)

struct __align__(8) cpx { float r, i; };      // 8-byte aligned: one 64-bit load or store per complex value

__host__ __device__ inline cpx c_mul(cpx a, cpx b) {
    cpx m;
    m.r = a.r * b.r - a.i * b.i;
    m.i = a.r * b.i + a.i * b.r;
    return m;
}
__host__ __device__ inline cpx c_add(cpx a, cpx b) { return cpx{a.r + b.r, a.i + b.i}; }
__host__ __device__ inline cpx c_sub(cpx a, cpx b) { return cpx{a.r - b.r, a.i - b.i}; }

constexpr int kSampleRate = 48000;
constexpr int kRnFrame = 480;      // RNNoise frame (Denoiser.zig:68-70)

inline size_t ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }

// Measurement / test knobs (kernel shapes, tile length, stream priorities ...): set ONLY through fvad_debug_set(); the
// library never reads the process environment, so a host's environment cannot change kernel selection.  Returns the
// value set for `key` (a copy valid until the calling thread's next query) or nullptr.
const char *debug_knob(const char *key);

// Device properties cached per device.
int device_sm_count(int device);
// Latency regime = few enough sequential chains that each can own an SM (biquad SM reservation, stage-stream priorities);
// otherwise throughput regime.  knob latency_mode=0/1 overrides the default rule (n_chains <= latency_chain_limit).
bool latency_regime(size_t n_chains, int device);

}  // namespace fvad
