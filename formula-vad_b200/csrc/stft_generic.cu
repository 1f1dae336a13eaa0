// stft_generic.cu -- FFT.fft / PipelineFFT.fft for every even n_fft other than the 2048 the pipeline defaults to
// (VAD.Config.fft_size and FFT.init(n_fft) accept any even size: src/AudioPipeline/VAD.zig:17-23, src/FFT.zig:28-58).
// k_stft2048 (stft.cu) is specialised to the radix-4^5 plan of the default size; this kernel executes KissFFT's plan for
// ANY complex length whose factors are 4, 2, 3, 5 -- kf_factor's order, 4s first (kiss_fft.c:307-328) -- stage by stage in
// shared memory, one CTA per frame, each butterfly by one thread in the reference's operation order
// (kf_bfly2 / kf_bfly3 / kf_bfly4 / kf_bfly5, kiss_fft.c:11-225), then kiss_fftr's split (kiss_fftr.c:63-117) and
// FFT.binOutput (FFT.zig:160-177).  It is a drop-in for correctness, not a bandwidth kernel: scalar f32 arithmetic, the
// plain recursion turned into an input permutation plus log-many in-place stages.
//
// kf_work (kiss_fft.c:235-300) decimates in time: level k (radix p_k, sub-length m_k, fstride_k = p_1 ... p_(k-1)) first
// runs its p_k sub-transforms on the inputs f[d + p_k j], then recombines blocks of p_k m_k outputs with kf_bfly.  Unrolled:
// input i = d_1 + p_1 (d_2 + p_2 (d_3 + ...)) lands at output position d_1 m_1 + d_2 m_2 + ... + d_L m_L, and the stages
// run from the innermost level outwards, level k applying kf_bfly(p_k) to every block of m_(k-1) = p_k m_k outputs.
#include "stft.cuh"
#include <cmath>
#include <mutex>
#include <vector>

namespace fvad {

namespace {

constexpr int kGenThreads = 128;

struct GenPlan {                       // by value into the kernel
    int n_fft, ncfft, nfac;
    int p[16], m[16];                  // kf_factor: p[k] * m[k] = m[k-1], m[-1] = ncfft
};

__device__ __forceinline__ cpx cmulf(cpx a, cpx b) {       // C_MUL, _kiss_fft_guts.h:87-89
    cpx r;
    r.r = a.r * b.r - a.i * b.i;
    r.i = a.r * b.i + a.i * b.r;
    return r;
}

__device__ __forceinline__ void bfly2(cpx *F, int m, const cpx *tw, int fs, int u) {                 // kiss_fft.c:11-36
    const cpx t = cmulf(F[m], tw[fs * u]);
    F[m].r = F[0].r - t.r; F[m].i = F[0].i - t.i;
    F[0].r += t.r; F[0].i += t.i;
}

__device__ __forceinline__ void bfly3(cpx *F, int m, const cpx *tw, int fs, int u) {                 // kiss_fft.c:86-129
    const int m2 = 2 * m;
    const cpx epi3 = tw[fs * m];
    cpx s1 = cmulf(F[m], tw[fs * u]), s2 = cmulf(F[m2], tw[2 * fs * u]);
    cpx s3{s1.r + s2.r, s1.i + s2.i}, s0{s1.r - s2.r, s1.i - s2.i};
    F[m].r = F[0].r - s3.r * 0.5f;                                   // HALF_OF(x) = x * (kiss_fft_scalar).5
    F[m].i = F[0].i - s3.i * 0.5f;
    s0.r *= epi3.i; s0.i *= epi3.i;                                  // C_MULBYSCALAR
    F[0].r += s3.r; F[0].i += s3.i;
    F[m2].r = F[m].r + s0.i;
    F[m2].i = F[m].i - s0.r;
    F[m].r -= s0.i;
    F[m].i += s0.r;
}

__device__ __forceinline__ void bfly4(cpx *F, int m, const cpx *tw, int fs, int u) {                 // kiss_fft.c:38-84, forward
    const int m2 = 2 * m, m3 = 3 * m;
    const cpx s0 = cmulf(F[m], tw[fs * u]), s1 = cmulf(F[m2], tw[2 * fs * u]), s2 = cmulf(F[m3], tw[3 * fs * u]);
    const cpx s5{F[0].r - s1.r, F[0].i - s1.i};
    F[0].r += s1.r; F[0].i += s1.i;
    const cpx s3{s0.r + s2.r, s0.i + s2.i}, s4{s0.r - s2.r, s0.i - s2.i};
    F[m2].r = F[0].r - s3.r; F[m2].i = F[0].i - s3.i;
    F[0].r += s3.r; F[0].i += s3.i;
    F[m].r = s5.r + s4.i;
    F[m].i = s5.i - s4.r;
    F[m3].r = s5.r - s4.i;
    F[m3].i = s5.i + s4.r;
}

__device__ __forceinline__ void bfly5(cpx *F, int m, const cpx *tw, int fs, int u) {                 // kiss_fft.c:131-193
    const cpx ya = tw[fs * m], yb = tw[fs * 2 * m];
    cpx *F0 = F, *F1 = F + m, *F2 = F + 2 * m, *F3 = F + 3 * m, *F4 = F + 4 * m;
    const cpx c0 = *F0;
    const cpx c1 = cmulf(*F1, tw[u * fs]), c2 = cmulf(*F2, tw[2 * u * fs]), c3 = cmulf(*F3, tw[3 * u * fs]), c4 = cmulf(*F4, tw[4 * u * fs]);
    const cpx c7{c1.r + c4.r, c1.i + c4.i}, c10{c1.r - c4.r, c1.i - c4.i}, c8{c2.r + c3.r, c2.i + c3.i}, c9{c2.r - c3.r, c2.i - c3.i};
    F0->r += c7.r + c8.r;
    F0->i += c7.i + c8.i;
    cpx c5, c6, c11, c12;
    c5.r = c0.r + c7.r * ya.r + c8.r * yb.r;
    c5.i = c0.i + c7.i * ya.r + c8.i * yb.r;
    c6.r = c10.i * ya.i + c9.i * yb.i;
    c6.i = -(c10.r * ya.i) - c9.r * yb.i;
    F1->r = c5.r - c6.r; F1->i = c5.i - c6.i;
    F4->r = c5.r + c6.r; F4->i = c5.i + c6.i;
    c11.r = c0.r + c7.r * yb.r + c8.r * ya.r;
    c11.i = c0.i + c7.i * yb.r + c8.i * ya.r;
    c12.r = -(c10.i * yb.i) + c9.i * ya.i;
    c12.i = c10.r * yb.i - c9.r * ya.i;
    F2->r = c11.r + c12.r; F2->i = c11.i + c12.i;
    F3->r = c11.r - c12.r; F3->i = c11.i - c12.i;
}

__global__ void __launch_bounds__(kGenThreads)
k_stft_generic(GenPlan P, const float *__restrict__ pcm, size_t n_rows, size_t row_stride, size_t n_frames, size_t hop,
               const float *__restrict__ window, const cpx *__restrict__ tw, const cpx *__restrict__ stw,
               const int *__restrict__ perm, float norm_factor, StftBands bands, float *__restrict__ d_band,
               size_t band_frame_stride, size_t band_frame_offset, float *__restrict__ d_spectra) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx *buf = reinterpret_cast<cpx *>(smem_raw);                   // ncfft
    float *mags = reinterpret_cast<float *>(buf + P.ncfft);         // ncfft + 1
    const int t = threadIdx.x, nc = P.ncfft, bins = nc + 1;
    const size_t total = n_rows * n_frames;
    for (size_t id = blockIdx.x; id < total; id += gridDim.x) {
        const size_t r = id / n_frames, f = id - r * n_frames;
        const float *x = pcm + r * row_stride + f * hop;
        // FFT.loadSamples (FFT.zig:147-158): sample * window; kiss_fftr reads the reals as ncfft complex values
        for (int i = t; i < nc; i += kGenThreads) {
            cpx v;
            v.r = x[2 * i] * window[2 * i];
            v.i = x[2 * i + 1] * window[2 * i + 1];
            buf[perm[i]] = v;
        }
        __syncthreads();
        for (int k = P.nfac - 1; k >= 0; k--) {
            const int p = P.p[k], m = P.m[k], fs = nc / (p * m), n_bf = nc / p;
            for (int j = t; j < n_bf; j += kGenThreads) {
                const int b = j / m, u = j - b * m;
                cpx *F = buf + (size_t)b * p * m + u;
                if (p == 4) bfly4(F, m, tw, fs, u);
                else if (p == 2) bfly2(F, m, tw, fs, u);
                else if (p == 3) bfly3(F, m, tw, fs, u);
                else bfly5(F, m, tw, fs, u);
            }
            __syncthreads();
        }
        // kiss_fftr.c:84-116 + FFT.binOutput (FFT.zig:160-177): magnitude = sqrt(r*r + i*i) * norm_factor
        for (int k = t; k <= nc / 2; k += kGenThreads) {
            if (k == 0) {
                const cpx d = buf[0];
                const float r0 = d.r + d.i, rn = d.r - d.i;
                const float z = 0.0f * 0.0f;
                mags[0] = sqrtf(r0 * r0 + z) * norm_factor;
                mags[nc] = sqrtf(rn * rn + z) * norm_factor;
            } else {
                const cpx fpk = buf[k];
                cpx fpnk = buf[nc - k];
                fpnk.i = -fpnk.i;
                const cpx f1k{fpk.r + fpnk.r, fpk.i + fpnk.i}, f2k{fpk.r - fpnk.r, fpk.i - fpnk.i};
                const cpx w = cmulf(f2k, stw[k - 1]);
                const float re = (f1k.r + w.r) * 0.5f, im = (f1k.i + w.i) * 0.5f;
                const float re2 = (f1k.r - w.r) * 0.5f, im2 = (w.i - f1k.i) * 0.5f;
                // when 2k == ncfft the reference writes freqdata[k] twice: the second assignment wins
                mags[nc - k] = sqrtf(re2 * re2 + im2 * im2) * norm_factor;
                if (2 * k != nc) mags[k] = sqrtf(re * re + im * im) * norm_factor;
            }
        }
        __syncthreads();
        if (t < bands.n_bands) {                                     // averageVolumeInBand: in-order f32 sum of bins lo..hi
            float v = 0.0f;
            for (int i = bands.lo[t]; i <= bands.hi[t]; i++) v += mags[i];
            d_band[((size_t)t * n_rows + r) * band_frame_stride + band_frame_offset + f] = v;
        }
        if (d_spectra != nullptr) {
            float *dst = d_spectra + id * (size_t)bins;
            for (int k = t; k < bins; k += kGenThreads) dst[k] = mags[k];
        }
        __syncthreads();
    }
}

struct GenTables {
    GenPlan plan{};
    float *d_window = nullptr;
    cpx *d_tw = nullptr, *d_stw = nullptr;
    int *d_perm = nullptr;
    float norm_factor = 0;
    size_t smem = 0;
};
std::mutex g_gen_mu;
std::vector<std::pair<std::pair<int, size_t>, GenTables>> g_gen;     // (device, n_fft) -> tables

int gen_tables_get(int device, size_t n_fft, const GenTables **out) {
    std::lock_guard<std::mutex> lk(g_gen_mu);
    for (auto &e : g_gen)
        if (e.first.first == device && e.first.second == n_fft) { *out = &e.second; return FVAD_OK; }
    GenTables t;
    const int nc = (int)(n_fft / 2);
    t.plan.n_fft = (int)n_fft; t.plan.ncfft = nc;
    {   // kf_factor (kiss_fft.c:307-328)
        int n = nc, p = 4, k = 0;
        const double floor_sqrt = floor(sqrt((double)n));
        do {
            while (n % p) {
                switch (p) { case 4: p = 2; break; case 2: p = 3; break; default: p += 2; break; }
                if (p > floor_sqrt) p = n;
            }
            n /= p;
            if (p > 5 || k >= 16)
                return set_error(FVAD_EUNSUPPORTED, "n_fft %zu: the complex length %d has the prime factor %d; this build covers "
                                 "lengths made of 2, 3, 4 and 5 (kf_bfly_generic is not implemented)", n_fft, nc, p);
            t.plan.p[k] = p; t.plan.m[k] = n; k++;
        } while (n > 1);
        t.plan.nfac = k;
    }
    std::vector<float> w(n_fft);
    host_hann_window_periodic(w.data(), n_fft);
    t.norm_factor = host_window_norm_factor(w.data(), n_fft) / (float)(n_fft / 2);                  // FFT.zig:94,165
    std::vector<cpx> tw(nc), stw(nc / 2 > 0 ? nc / 2 : 1);
    for (int i = 0; i < nc; i++) {                                   // kiss_fft.c:357-363
        const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
        const double phase = -2 * pi * i / nc;
        tw[i].r = (float)cos(phase); tw[i].i = (float)sin(phase);
    }
    for (int i = 0; i < nc / 2; i++) {                               // kiss_fftr.c:53-59
        const double phase = -3.14159265358979323846264338327 * ((double)(i + 1) / nc + .5);
        stw[i].r = (float)cos(phase); stw[i].i = (float)sin(phase);
    }
    std::vector<int> perm(nc);
    for (int i = 0; i < nc; i++) {                                   // digit d_k of i (radix p_k) goes to weight m_k
        int rest = i, o = 0;
        for (int k = 0; k < t.plan.nfac; k++) { o += (rest % t.plan.p[k]) * t.plan.m[k]; rest /= t.plan.p[k]; }
        perm[i] = o;
    }
    t.smem = sizeof(cpx) * nc + sizeof(float) * (nc + 1) + 16;
    int prev = 0;
    FVAD_CUDA_TRY(cudaGetDevice(&prev));
    FVAD_CUDA_TRY(cudaSetDevice(device));
    cudaError_t e = cudaMalloc(&t.d_window, sizeof(float) * n_fft);
    if (e == cudaSuccess) e = cudaMalloc(&t.d_tw, sizeof(cpx) * tw.size());
    if (e == cudaSuccess) e = cudaMalloc(&t.d_stw, sizeof(cpx) * stw.size());
    if (e == cudaSuccess) e = cudaMalloc(&t.d_perm, sizeof(int) * perm.size());
    if (e == cudaSuccess) e = cudaMemcpy(t.d_window, w.data(), sizeof(float) * n_fft, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t.d_tw, tw.data(), sizeof(cpx) * tw.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t.d_stw, stw.data(), sizeof(cpx) * stw.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t.d_perm, perm.data(), sizeof(int) * perm.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && t.smem > 48 * 1024)
        e = cudaFuncSetAttribute(k_stft_generic, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaSetDevice(prev);
    if (e != cudaSuccess) return set_error(FVAD_ECUDA, "STFT tables for n_fft %zu: %s (no CPU fallback exists)", n_fft, cudaGetErrorString(e));
    g_gen.push_back({{device, n_fft}, t});
    *out = &g_gen.back().second;
    return FVAD_OK;
}

}  // namespace

int stft_generic_supported(size_t n_fft) {
    if (n_fft < 4 || (n_fft & 1)) return set_error(FVAD_EINVALID_FFT_SIZE, "n_fft %zu must be even and at least 4", n_fft);
    if (n_fft > 32768) return set_error(FVAD_EUNSUPPORTED, "n_fft %zu: larger than the 32768 this build's shared-memory plan holds", n_fft);
    size_t n = n_fft / 2;
    for (size_t p : {(size_t)2, (size_t)3, (size_t)5}) while (n % p == 0) n /= p;
    if (n != 1)
        return set_error(FVAD_EUNSUPPORTED, "n_fft %zu: the complex length %zu has a prime factor above 5 (kf_bfly_generic is not "
                         "implemented in this build)", n_fft, n_fft / 2);
    return FVAD_OK;
}

int stft_generic_norm_factor(int device, size_t n_fft, float *norm) {
    const GenTables *t = nullptr;
    FVAD_TRY(gen_tables_get(device, n_fft, &t));
    *norm = t->norm_factor;
    return FVAD_OK;
}

int stft_generic_launch(int device, size_t n_fft, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames, size_t hop,
                        const StftBands &bands, float *d_band, size_t band_frame_stride, size_t band_frame_offset, float *d_spectra,
                        const float *d_window_override, float norm_override, cudaStream_t stream) {
    if (n_rows == 0 || n_frames == 0) return FVAD_OK;
    FVAD_TRY(stft_generic_supported(n_fft));
    const GenTables *t = nullptr;
    FVAD_TRY(gen_tables_get(device, n_fft, &t));
    for (int i = 0; i < bands.n_bands; i++)
        if (bands.lo[i] < 0 || bands.hi[i] > (int)(n_fft / 2) || bands.hi[i] < bands.lo[i])
            return set_error(FVAD_EOUT_OF_RANGE, "band %d..%d outside the %zu bins", bands.lo[i], bands.hi[i], n_fft / 2 + 1);
    const size_t total = n_rows * n_frames;
    const size_t grid = std::min<size_t>(total, (size_t)device_sm_count(device) * 8);
    FVAD_LAUNCH(k_stft_generic, (unsigned)grid, kGenThreads, t->smem, stream, t->plan, d_pcm, n_rows, row_stride, n_frames, hop,
                d_window_override ? d_window_override : t->d_window, t->d_tw, t->d_stw, t->d_perm,
                d_window_override ? norm_override : t->norm_factor, bands, d_band, band_frame_stride, band_frame_offset, d_spectra);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad
