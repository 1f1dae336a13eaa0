// noalloc.cu -- fvad_vadm_sweep_static: the VADMachine sweep over shared features (VAD.zig:171-183,375-380;
// BASELINE config 5) for callers that may not allocate: code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  It has no counterpart in the
// reference.
//
// The scans' scratch -- every machine's three f64 rings (RollingAverage.zig:5-9) and the derived parameters --
// lives in ONE file-scope device array of FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES, held by every GPU that loads the
// library until the process exits.  Inputs and outputs are the caller's device buffers.  The entry point calls no
// cudaMalloc / cudaFree / allocating library routine; requests that do not fit the array are rejected.
// Every call shares the array and finds the previous call's contents in it: the kernel clears the ring columns
// it is going to use before the first record.  Calls that overlap on the GPU (two streams, two host threads)
// race on the array -- serialise them.
//
// The kernel is the sweep's (vadm_sweep.cu: VADMachine.run, src/AudioPipeline/VADMachine.zig:126-310, one device thread
// per (stream, config) walking all records of one launch with its state in registers); this file only supplies the
// scratch.
#include "common.cuh"
#include "vadm.cuh"
#include <algorithm>
#include <mutex>

using namespace fvad;

namespace {

constexpr size_t kScratchBytes = FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES;
constexpr int kMaxConfigs = FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS;
constexpr size_t kParamBytes = ((size_t)kMaxConfigs * sizeof(VadmParams) + 255) & ~(size_t)255;
static_assert(kParamBytes < kScratchBytes, "scratch too small for the parameter block");

// In HOST code this name is only a stand-in for the symbol, never a device address: the address a kernel may
// use comes from cudaGetSymbolAddress, once per device (scratch_for_current_device).
static __device__ __align__(256) unsigned char g_sweep_scratch[kScratchBytes];

std::mutex g_scratch_mu;
void *g_scratch_ptr[64] = {};

int scratch_for_current_device(unsigned char **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return set_error(FVAD_EUNSUPPORTED, "device ordinal %d out of the supported range", dev);
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sweep_scratch);          // with device `dev` current: its copy of the array
        if (e != cudaSuccess || !p)
            return set_error(FVAD_ECUDA, "cudaGetSymbolAddress(g_sweep_scratch) on device %d: %s (no CPU fallback exists)", dev,
                             cudaGetErrorString(e));
        g_scratch_ptr[dev] = p;
    }
    *out = (unsigned char *)g_scratch_ptr[dev];
    return FVAD_OK;
}

}  // namespace

extern "C" int fvad_vadm_sweep_static(const float *d_min_volume, const float *d_vol_ratio, const float *d_rnn_vad,
                                      size_t n_streams, size_t n_records, uint64_t first_index,
                                      const fvad_vad_machine_config *configs, size_t n_configs, size_t sample_rate,
                                      size_t fft_size, fvad_segment *d_segs, size_t max_segs, uint32_t *d_counts,
                                      void *cuda_stream) {
    if (!d_min_volume || !d_vol_ratio || !configs || !d_segs || !d_counts || !n_streams || !n_configs || !n_records || !max_segs)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_vadm_sweep_static: null/zero argument");
    if (fft_size == 0) return set_error(FVAD_EINVALID_FFT_SIZE, "fft_size 0");
    if (n_configs > (size_t)kMaxConfigs)
        return set_error(FVAD_EOUT_OF_MEMORY, "fvad_vadm_sweep_static: %zu configs exceed the static limit of %d "
                         "(FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS); use fvad_vadm_sweep", n_configs, kMaxConfigs);
    VadmParams hp[kMaxConfigs];
    size_t max_long = 0, max_short = 0, max_ratio = 0;
    for (size_t k = 0; k < n_configs; k++) {
        size_t nd = 0;
        FVAD_TRY(vadm_params_from_config(configs[k], sample_rate, fft_size, &hp[k], &nd));
        max_long = std::max<size_t>(max_long, hp[k].long_len);
        max_short = std::max<size_t>(max_short, hp[k].short_len);
        max_ratio = std::max<size_t>(max_ratio, hp[k].ratio_len);
    }
    for (size_t k = 0; k < n_configs; k++) hp[k].band = 0;      // one min_volume plane
    const size_t rows = max_long + max_short + max_ratio;
    const size_t n_scans = n_streams * n_configs;
    if (n_scans > (size_t)1 << 24 || rows > (kScratchBytes - kParamBytes) / sizeof(double) / n_scans)
        return set_error(FVAD_EOUT_OF_MEMORY, "fvad_vadm_sweep_static: %zu scans x %zu ring entries x 8 B = %.1f MiB exceed the "
                         "static scratch of %.1f MiB (FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES); use fvad_vadm_sweep or fewer "
                         "streams / configs per call", n_scans, rows, (double)n_scans * rows * 8 / 1048576.0,
                         (double)(kScratchBytes - kParamBytes) / 1048576.0);
    unsigned char *scratch = nullptr;
    FVAD_TRY(scratch_for_current_device(&scratch));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // `hp` is pageable: the runtime has staged it by the time this returns, so the stack array may go out of scope
    FVAD_CUDA_TRY(cudaMemcpyAsync(scratch, hp, sizeof(VadmParams) * n_configs, cudaMemcpyHostToDevice, st));
    VadmSweepIo io{};
    io.min_volume = d_min_volume; io.band_stride = 0; io.vol_ratio = d_vol_ratio; io.vad = d_rnn_vad;
    io.rec_stride = n_records; io.n_records = n_records;
    io.first_index = first_index; io.index_step = fft_size;
    io.segs = d_segs; io.max_segs = max_segs; io.counts = d_counts; io.fallbacks = nullptr;
    FVAD_TRY(vadm_sweep_launch((const VadmParams *)scratch, (double *)(scratch + kParamBytes), (int)max_long, (int)max_short,
                               n_streams, (int)n_configs, io, st));
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}
