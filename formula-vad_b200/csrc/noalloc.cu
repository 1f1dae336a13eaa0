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
// The machine itself is VADMachine.run (src/AudioPipeline/VADMachine.zig:126-310) over RollingAverage
// (src/structures/RollingAverage.zig:34-56), one device thread per (stream, config) walking all records of one
// launch with its state in registers.  Ring entry i of scan `id` lives at rings[i * n_scans + id], so that the 32
// scans of a warp touch one 256-byte line per ring index.
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

// ---- rings, column layout -------------------------------------------------------------------------------------
struct Col {                       // one scan's view of a ring: entry i at p[i * pitch]
    double *p;
    size_t pitch;
    __device__ __forceinline__ double &operator[](int i) const { return p[(size_t)i * pitch]; }
};

__device__ __forceinline__ double col_avg_exact(const Col &r, int cnt) {               // RollingAverage.zig:45-56
    double a = 0.0;
    const double scalar = 1.0 / (double)cnt;
    for (int i = 0; i < cnt; i++) a += r[i] * scalar;
    return a;
}

__device__ __forceinline__ double col_push_exact(const Col &r, int len, int &w, int &cnt, float sample) {
    r[w] = (double)sample;                                                              // RollingAverage.zig:34-43
    if (++w == len) w = 0;
    if (cnt < len) cnt++;
    return col_avg_exact(r, cnt);
}

__device__ __forceinline__ double col_sum_blocked(const Col &r, int cnt) {             // raw sum, four partial sums
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    int i = 0;
    for (; i + 3 < cnt; i += 4) { a0 += r[i]; a1 += r[i + 1]; a2 += r[i + 2]; a3 += r[i + 3]; }
    for (; i < cnt; i++) a0 += r[i];
    return (a0 + a1) + (a2 + a3);
}

// The long-term ring (4218 entries by default) is re-summed in index order on every push in the reference; its
// value feeds one comparison, `short_term > last_avg * factor` (VADMachine.zig:158-160).  This kernel carries a
// running RAW sum `rsum` of the ring with a running bound `err` on |rsum - (real sum of the ring)|:
//   full ring   rsum' = fl(rsum + fl(nv - old))   err += 2^-52 (|rsum'| + |nv| + |old|)
//   filling     rsum' = fl(rsum + nv)             err += 2^-52 |rsum'|
//   refresh     rsum  = four-way blocked sum      err  = rsum * (n/4 + 4) * 2^-52
// (every term is twice the rounding error it covers).  With all entries >= 0 the reference's own value is
// R = fl(1/n) * realsum * (1 + e), |e| <= n * 2^-53, so |fl(fl(rsum * fl(1/n)) * factor) - fl(R * factor)| is at most
// thr * k_rel + err * fl(1/n) * factor * (1 + 1e-6) with k_rel = 1e-9 + n * 2^-52.  Outside that margin the comparison
// is decided by rsum; inside it -- or if a negative / NaN value ever entered the ring -- the reference's in-order
// sum is evaluated.  The running sum is refreshed every 512 pushes and whenever err exceeds 2^-30 * rsum (a large
// entry leaving the ring cancels the sum while the bound keeps its size).  The decision is the reference's always.
struct LongRing {
    int w, cnt, since;
    bool has_last, exact, poisoned;
    double last, rsum, err, scalar;       // scalar = 1 / cnt
};

__device__ __forceinline__ void long_refresh(const Col &r, LongRing &L) {
    L.rsum = col_sum_blocked(r, L.cnt);
    L.err = L.rsum * ((double)(L.cnt / 4 + 4) * 0x1p-52);
    L.since = 0;
}

__device__ __forceinline__ void long_push(const Col &r, int len, LongRing &L, float sample) {
    const double nv = (double)sample;
    if (!(nv >= 0.0)) L.poisoned = true;
    if (L.cnt < len) {
        r[L.w] = nv;
        L.cnt++;
        L.scalar = 1.0 / (double)L.cnt;
        L.rsum += nv;
        L.err += 0x1p-52 * fabs(L.rsum);
    } else {
        const double old = r[L.w];
        r[L.w] = nv;
        L.rsum += (nv - old);
        L.err += 0x1p-52 * (fabs(L.rsum) + fabs(nv) + fabs(old));
    }
    if (++L.w == len) L.w = 0;
    L.has_last = true; L.exact = false;
    if (++L.since >= 512 || !(L.err <= L.rsum * 0x1p-30)) long_refresh(r, L);
}

__global__ void __launch_bounds__(32)
k_vadm_sweep_static(const VadmParams *__restrict__ params, double *__restrict__ rings, int max_long, int max_short,
                    const float *__restrict__ minv, const float *__restrict__ ratio, const float *__restrict__ vad,
                    size_t n_streams, int n_configs, size_t n_records, unsigned long long first_index,
                    unsigned long long index_step, fvad_segment *__restrict__ segs, size_t max_segs,
                    uint32_t *__restrict__ counts) {
    const size_t n_scans = n_streams * (size_t)n_configs;
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_scans) return;
    const size_t s = id / n_configs;
    const VadmParams P = params[id - s * n_configs];
    const Col ring_long{rings + id, n_scans};
    const Col ring_short{rings + (size_t)max_long * n_scans + id, n_scans};
    const Col ring_ratio{rings + ((size_t)max_long + max_short) * n_scans + id, n_scans};

    // the array keeps the previous call's contents: every column used below is rewritten first
    for (int i = 0; i < P.long_len; i++) ring_long[i] = P.has_initial ? P.initial : 0.0;   // RollingAverage.zig:19-25
    for (int i = 0; i < P.short_len; i++) ring_short[i] = 0.0;
    for (int i = 0; i < P.ratio_len; i++) ring_ratio[i] = 0.0;
    LongRing L{};
    if (P.has_initial) {
        L.cnt = P.long_len;
        L.scalar = 1.0 / (double)L.cnt;
        L.last = col_avg_exact(ring_long, L.cnt);
        L.has_last = true; L.exact = true;
        L.poisoned = !(P.initial >= 0.0);
        long_refresh(ring_long, L);
    }
    int state = 0, short_w = 0, short_cnt = 0, ratio_w = 0, ratio_cnt = 0;
    unsigned long long start_idx = 0, end_idx = 0;
    float sp_vad = 0.f, sp_ratio = 0.f;
    unsigned sp_vad_cnt = 0, sp_ratio_cnt = 0;
    uint32_t n_out = 0;
    const double k_rel = 1e-9 + (double)P.long_len * 0x1p-52;
    const float *mv = minv + s * n_records, *rt = ratio + s * n_records, *vd = vad ? vad + s * n_records : nullptr;

    constexpr int kChunk = 8;
    for (size_t f0 = 0; f0 < n_records; f0 += kChunk) {
        float c_mv[kChunk], c_rt[kChunk], c_vd[kChunk];
#pragma unroll
        for (int j = 0; j < kChunk; j++) {                       // the chunk's records: independent loads, issued together
            const size_t f = f0 + j < n_records ? f0 + j : n_records - 1;
            c_mv[j] = __ldg(mv + f); c_rt[j] = __ldg(rt + f); c_vd[j] = vd ? __ldg(vd + f) : 0.0f;
        }
#pragma unroll 1
        for (int j = 0; j < kChunk && f0 + j < n_records; j++) {
            const float min_volume = c_mv[j], volume_ratio = c_rt[j], v = c_vd[j];
            const unsigned long long index = first_index + (f0 + j) * index_step;

            const double short_term = col_push_exact(ring_short, P.short_len, short_w, short_cnt, min_volume);  // VADMachine.zig:155
            const double cvr = col_push_exact(ring_ratio, P.ratio_len, ratio_w, ratio_cnt, volume_ratio);       // :156

            bool primary;
            if (!L.has_last) {
                const double base = P.has_initial ? P.initial : short_term;                                       // :158
                primary = short_term > base * P.factor;
            } else {
                bool decided = false;
                if (!L.exact && !L.poisoned) {
                    const double scalar = L.scalar;
                    const double thr = (L.rsum * scalar) * P.factor;
                    if (L.err == 0.0) { primary = short_term > thr; decided = true; }      // every entry is zero: rsum is the sum
                    else {
                        const double tol = thr * k_rel + (L.err * scalar) * P.factor * 1.000001;
                        if (short_term > thr + tol) { primary = true; decided = true; }
                        else if (short_term < thr - tol) { primary = false; decided = true; }
                    }
                }
                if (!decided) {
                    if (!L.exact) { L.last = col_avg_exact(ring_long, L.cnt); L.exact = true; }   // the reference's own sum
                    primary = short_term > L.last * P.factor;                                     // :159-160
                }
            }
            const bool met = primary && (cvr > P.ratio_threshold);
            if (!met) long_push(ring_long, P.long_len, L, min_volume);                            // :165-167

            const int from_state = state;
            switch (state) {                                                                      // :176-221
            case 0:
                if (met) { state = 1; start_idx = index; }
                break;
            case 1:
                if (met && index - start_idx >= P.min_consecutive) state = 2;
                else if (!met) state = 0;
                break;
            case 2:
                if (!met) { state = 3; end_idx = index; }
                break;
            case 3:
                if (met) state = 2;
                else if (index - end_idx >= P.max_gap) {
                    state = 0;
                    // onSpeechEnd, VADMachine.zig:252-294
                    const float length_realtime = (float)(end_idx - start_idx) / P.sample_rate_f;
                    if (length_realtime >= P.min_dur) {
                        if (n_out < max_segs) {
                            fvad_segment seg;
                            seg.sample_from = P.start_buffer > start_idx ? 0 : start_idx - P.start_buffer;
                            seg.sample_to = end_idx + P.end_buffer;
                            seg.debug_rnn_vad = sp_vad / (float)sp_vad_cnt;
                            seg.debug_avg_speech_vol_ratio = sp_ratio / (float)sp_ratio_cnt;
                            segs[id * max_segs + n_out] = seg;
                        }
                        n_out++;
                    }
                }
                break;
            }
            if (from_state == 0 && state == 1) {                                                  // trackSpeechStats, :232-250
                sp_vad = v; sp_vad_cnt = 1; sp_ratio = volume_ratio; sp_ratio_cnt = 1;
            } else if (from_state == 1 || from_state == 2) {
                sp_vad += v; sp_vad_cnt += 1; sp_ratio += volume_ratio; sp_ratio_cnt += 1;
            }
        }
    }
    counts[id] = n_out;
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
    FVAD_LAUNCH(k_vadm_sweep_static, (unsigned)ceil_div(n_scans, 32), 32, 0, st, (const VadmParams *)scratch,
                (double *)(scratch + kParamBytes), (int)max_long, (int)max_short, d_min_volume, d_vol_ratio, d_rnn_vad,
                n_streams, (int)n_configs, n_records, (unsigned long long)first_index, (unsigned long long)fft_size, d_segs,
                max_segs, d_counts);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}
