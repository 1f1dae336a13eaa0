// stft.cu -- see stft.cuh.  One CTA of 64 threads transforms one 2048-sample frame at a
// time and loops over frames (persistent grid sized to the SM count).
//
// Arithmetic follows KissFFT's float build operation for operation (radix-4 DIT,
// 1024 complex points = five radix-4 stages m=1,4,16,64,256, then the kiss_fftr split
// pass), so spectra are bit-identical to the reference's.  What is re-organised is only
// WHERE values live: stages (m=1,m=4) and (m=16,m=64) are fused in registers (16 points
// per thread), the last stage + split pass + magnitude are fused and, when only the
// speech band is wanted, pruned to the bins that are read.  The only arithmetic that is
// skipped is multiplication by the twiddle W^0 = (1,-0): x*(1,-0) == x for every finite x
// up to the sign of a zero, which no magnitude can observe.
#include "stft.cuh"
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace fvad {

// ------------------------------------------------------------------------------------------
// host tables
// ------------------------------------------------------------------------------------------
// hannWindowPeriodic -> cosineSumWindowPeriodic (window_fn.zig:22-28,51-68): K=1, alphas
// {0.5,0.5}, evaluated in f32; `@cos` on f32 is taken as the correctly rounded cosine.
void host_hann_window_periodic(float *w, size_t n) {
    const float N = (float)n;
    const float two_pi = (float)(2 * 3.14159265358979323846264338327950288419716939937510);
    for (size_t i = 0; i < n; i++) {
        float nf = (float)i;
        float r = 0;
        r += (1.0f * 0.5f) * (float)cos((double)(((two_pi * 0.0f) * nf) / N));
        r += (-1.0f * 0.5f) * (float)cos((double)(((two_pi * 1.0f) * nf) / N));
        w[i] = r;
    }
}
float host_window_norm_factor(const float *w, size_t n) {   // window_fn.zig:8-16
    float sum = 0;
    for (size_t i = 0; i < n; i++) sum += w[i];
    return (float)n / sum;
}

static std::mutex g_tab_mu;
static StftTables g_tabs[64];
static bool g_tab_ready[64];

int stft_tables_get(int device, const StftTables **out) {
    if (device < 0 || device >= 64) return set_error(FVAD_EINVALID_ARGUMENT, "device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_tab_mu);
    if (!g_tab_ready[device]) {
        StftTables t;
        std::vector<float> w(kFft);
        host_hann_window_periodic(w.data(), kFft);
        float wn = host_window_norm_factor(w.data(), kFft);
        t.norm_factor = wn / (float)(kFft / 2);
        std::vector<cpx> tw(kFftC), stw(kFftC / 2);
        for (int i = 0; i < kFftC; i++) {               // kiss_fft.c:357-363
            const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
            double phase = -2 * pi * i / kFftC;
            tw[i].r = (float)cos(phase);
            tw[i].i = (float)sin(phase);
        }
        for (int i = 0; i < kFftC / 2; i++) {           // kiss_fftr.c:53-59
            double phase = -3.14159265358979323846264338327 * ((double)(i + 1) / kFftC + .5);
            stw[i].r = (float)cos(phase);
            stw[i].i = (float)sin(phase);
        }
        int prev = 0;
        FVAD_CUDA_TRY(cudaGetDevice(&prev));
        FVAD_CUDA_TRY(cudaSetDevice(device));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_window, sizeof(float) * kFft));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_tw, sizeof(cpx) * kFftC));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_stw, sizeof(cpx) * kFftC / 2));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_twB, sizeof(cpx) * 15 * 64));
        {   // pass-B twiddles per thread t = (pb, jj): row-major [15][64] so that a warp reads 256 contiguous bytes
            std::vector<cpx> twb(15 * 64);
            for (int th = 0; th < 64; th++) {
                const int jj = th & 15;
                for (int q = 1; q < 4; q++) twb[(q - 1) * 64 + th] = tw[jj * 16 * q];                       // stage m=16
                for (int g = 0; g < 4; g++)
                    for (int q = 1; q < 4; q++) twb[(3 + 3 * g + q - 1) * 64 + th] = tw[(jj + 16 * g) * 4 * q];   // stage m=64
            }
            FVAD_CUDA_TRY(cudaMemcpy(t.d_twB, twb.data(), sizeof(cpx) * twb.size(), cudaMemcpyHostToDevice));
            for (int k = 1; k < 4; k++)
                for (int q = 1; q < 4; q++) t.twA.v[(k - 1) * 3 + q - 1] = tw[k * 64 * q];                 // stage m=4, uniform
        }
        FVAD_CUDA_TRY(cudaMemcpy(t.d_window, w.data(), sizeof(float) * kFft, cudaMemcpyHostToDevice));
        FVAD_CUDA_TRY(cudaMemcpy(t.d_tw, tw.data(), sizeof(cpx) * kFftC, cudaMemcpyHostToDevice));
        FVAD_CUDA_TRY(cudaMemcpy(t.d_stw, stw.data(), sizeof(cpx) * kFftC / 2, cudaMemcpyHostToDevice));
        FVAD_CUDA_TRY(cudaSetDevice(prev));
        g_tabs[device] = t;
        g_tab_ready[device] = true;
    }
    *out = &g_tabs[device];
    return FVAD_OK;
}

// ------------------------------------------------------------------------------------------
// device
// ------------------------------------------------------------------------------------------
// radix-4 butterfly, forward branch of kf_bfly4 (kiss_fft.c:38-84); s0,s1,s2 are the
// already twiddled inputs F[m]*tw1, F[2m]*tw2, F[3m]*tw3.
__device__ __forceinline__ void bfly4(cpx &f0, cpx &f1, cpx &f2, cpx &f3, cpx s0, cpx s1, cpx s2) {
    cpx s5 = c_sub(f0, s1);
    f0 = c_add(f0, s1);
    cpx s3 = c_add(s0, s2);
    cpx s4 = c_sub(s0, s2);
    f2 = c_sub(f0, s3);
    f0 = c_add(f0, s3);
    f1.r = s5.r + s4.i;
    f1.i = s5.i - s4.r;
    f3.r = s5.r - s4.i;
    f3.i = s5.i + s4.r;
}

__device__ __forceinline__ cpx ldg_cpx(const cpx *p) {        // 8-byte read-only global load
    const float2 v = __ldg(reinterpret_cast<const float2 *>(p));
    return cpx{v.x, v.y};
}

// shared-memory slot of complex position pos (0..1023): one float2 of padding per 16 and per
// 256 positions keeps all three access patterns (16-blocks, stride-16, stride-256) conflict-free.
__device__ __forceinline__ int slot(int pos) { return pos + (pos >> 4) + (pos >> 8); }
constexpr int kSlots = 1024 + 64 + 4 + 4;

template <bool FULL>
__global__ void __launch_bounds__(64, 16) k_stft2048(const float *__restrict__ pcm, size_t n_rows, size_t row_stride,
                                                 size_t n_frames, size_t hop, const float *__restrict__ window,
                                                 const cpx *__restrict__ tw, const cpx *__restrict__ stw,
                                                 const cpx *__restrict__ twB, const StftTwA twA,
                                                 float norm_factor, StftBands bands, float *__restrict__ d_band,
                                                 size_t band_frame_stride, size_t band_frame_offset,
                                                 float *__restrict__ d_spectra) {
    __shared__ float2 sm[kSlots];
    __shared__ float mags[FULL ? kBins + 3 : 256];
    const int t = threadIdx.x;

    // Twiddles are NOT kept in registers across frames (48 registers would halve the occupancy): pass A's are
    // uniform and arrive as kernel parameters (constant-bank operands), pass B's come from a compact
    // per-thread table that stays in L1.
    const int pb = t >> 4, jj = t & 15;
    // pass A block: thread t owns input residue n_low = t; its 16 points land in block blk
    const int blk = ((t & 3) << 4) | (t & 12) | (t >> 4);
    const float2 *win2 = reinterpret_cast<const float2 *>(window);

    const size_t total = n_rows * n_frames;
    for (size_t fid = blockIdx.x; fid < total; fid += gridDim.x) {
        const size_t r = fid / n_frames, f = fid - r * n_frames;
        const float2 *src = reinterpret_cast<const float2 *>(pcm + r * row_stride + f * hop);

        // ---------------- pass A: load + window, stages m=1 and m=4 in registers ------------
        cpx a[16];
#pragma unroll
        for (int e = 0; e < 16; e++) {
            float2 v = src[64 * e + t];                 // coalesced: 32 lanes x 8 B contiguous
            float2 w = win2[64 * e + t];
            // input n = t + 64 e = t + 64 (k4 + 4 k5) -> local slot i = 4 k4 + k5
            const int i = 4 * (e & 3) + (e >> 2);
            a[i].r = v.x * w.x;                         // FFT.zig:150,155 sample * window[idx]
            a[i].i = v.y * w.y;
        }
#pragma unroll
        for (int g = 0; g < 4; g++)                     // stage m=1: twiddles are all W^0
            bfly4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3]);
        bfly4(a[0], a[4], a[8], a[12], a[4], a[8], a[12]);   // stage m=4, k=0
#pragma unroll
        for (int k = 1; k < 4; k++)
            bfly4(a[k], a[k + 4], a[k + 8], a[k + 12], c_mul(a[k + 4], twA.v[(k - 1) * 3]), c_mul(a[k + 8], twA.v[(k - 1) * 3 + 1]),
                  c_mul(a[k + 12], twA.v[(k - 1) * 3 + 2]));
#pragma unroll
        for (int i = 0; i < 16; i++) sm[slot(16 * blk + i)] = make_float2(a[i].r, a[i].i);
        __syncthreads();

        // ---------------- pass B: stages m=16 and m=64 -------------------------------------
        const cpx *twb = twB + t;
        asm volatile("" : "+l"(twb));                   // opaque per frame: keeps the 15 loads inside the loop
        cpx twB1[3];
#pragma unroll
        for (int q = 0; q < 3; q++) twB1[q] = ldg_cpx(twb + q * 64);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            float2 v = sm[slot(256 * pb + jj + 16 * i)];
            a[i].r = v.x; a[i].i = v.y;
        }
#pragma unroll
        for (int g = 0; g < 4; g++) {                   // m=16: points jj+16*(4g+r), twiddle index jj
            // (lanes with jj == 0 multiply by W^0 = (1,-0) exactly as KissFFT does: no divergent special case)
            bfly4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3], c_mul(a[4 * g + 1], twB1[0]),
                  c_mul(a[4 * g + 2], twB1[1]), c_mul(a[4 * g + 3], twB1[2]));
        }
#pragma unroll
        for (int g = 0; g < 4; g++) {                   // m=64: points (jj+16g)+64r, twiddle index jj+16g
            const cpx w1 = ldg_cpx(twb + (3 + 3 * g) * 64), w2 = ldg_cpx(twb + (4 + 3 * g) * 64), w3 = ldg_cpx(twb + (5 + 3 * g) * 64);
            bfly4(a[g], a[g + 4], a[g + 8], a[g + 12], c_mul(a[g + 4], w1), c_mul(a[g + 8], w2), c_mul(a[g + 12], w3));
        }
#pragma unroll
        for (int i = 0; i < 16; i++) sm[slot(256 * pb + jj + 16 * i)] = make_float2(a[i].r, a[i].i);
        __syncthreads();

        // ---------------- pass C: stage m=256 + kiss_fftr split + |X|*norm -----------------
        if (FULL) {
#pragma unroll
            for (int it = 0; it < 4; it++) {
                const int k = t + 64 * it;
                float2 v0 = sm[slot(k)], v1 = sm[slot(k + 256)], v2 = sm[slot(k + 512)], v3 = sm[slot(k + 768)];
                cpx f0{v0.x, v0.y}, f1{v1.x, v1.y}, f2{v2.x, v2.y}, f3{v3.x, v3.y};
                if (k == 0) bfly4(f0, f1, f2, f3, f1, f2, f3);
                else bfly4(f0, f1, f2, f3, c_mul(f1, tw[k]), c_mul(f2, tw[2 * k]), c_mul(f3, tw[3 * k]));
                sm[slot(k)] = make_float2(f0.r, f0.i);
                sm[slot(k + 256)] = make_float2(f1.r, f1.i);
                sm[slot(k + 512)] = make_float2(f2.r, f2.i);
                sm[slot(k + 768)] = make_float2(f3.r, f3.i);
            }
            __syncthreads();
            for (int k = t; k <= 512; k += 64) {        // kiss_fftr.c:90-116
                if (k == 0) {
                    float2 d = sm[slot(0)];
                    float r0 = d.x + d.y, rn = d.x - d.y;
                    mags[0] = sqrtf(r0 * r0 + 0.0f * 0.0f) * norm_factor;
                    mags[1024] = sqrtf(rn * rn + 0.0f * 0.0f) * norm_factor;
                } else {
                    float2 p = sm[slot(k)], q = sm[slot(1024 - k)];
                    cpx fpk{p.x, p.y}, fpnk{q.x, -q.y};
                    cpx f1k = c_add(fpk, fpnk), f2k = c_sub(fpk, fpnk);
                    cpx w = c_mul(f2k, stw[k - 1]);
                    float re = (f1k.r + w.r) * 0.5f, im = (f1k.i + w.i) * 0.5f;
                    float re2 = (f1k.r - w.r) * 0.5f, im2 = (w.i - f1k.i) * 0.5f;
                    // k == 512: the reference writes freqdata[512] twice, the second wins
                    mags[1024 - k] = sqrtf(re2 * re2 + im2 * im2) * norm_factor;
                    if (k != 512) mags[k] = sqrtf(re * re + im * im) * norm_factor;   // FFT.zig:167-176
                }
            }
        } else {
            // only the bins some band reads (all within 1..255)
            const int lo = bands.lo[0], hi = bands.hi[bands.n_bands - 1];   // host guarantees a covering range
            for (int k = lo + t; k <= hi; k += 64) {
                cpx x0, y3;
                {   // output 0 of butterfly k = tmp[k]
                    float2 v0 = sm[slot(k)], v1 = sm[slot(k + 256)], v2 = sm[slot(k + 512)], v3 = sm[slot(k + 768)];
                    cpx f0{v0.x, v0.y};
                    cpx s0 = c_mul(cpx{v1.x, v1.y}, tw[k]), s1 = c_mul(cpx{v2.x, v2.y}, tw[2 * k]),
                        s2 = c_mul(cpx{v3.x, v3.y}, tw[3 * k]);
                    f0 = c_add(f0, s1);
                    x0 = c_add(f0, c_add(s0, s2));
                }
                {   // output 3 of butterfly 256-k = tmp[1024-k]
                    const int kk = 256 - k;
                    float2 v0 = sm[slot(kk)], v1 = sm[slot(kk + 256)], v2 = sm[slot(kk + 512)], v3 = sm[slot(kk + 768)];
                    cpx f0{v0.x, v0.y};
                    cpx s0 = c_mul(cpx{v1.x, v1.y}, tw[kk]), s1 = c_mul(cpx{v2.x, v2.y}, tw[2 * kk]),
                        s2 = c_mul(cpx{v3.x, v3.y}, tw[3 * kk]);
                    cpx s5 = c_sub(f0, s1), s4 = c_sub(s0, s2);
                    y3.r = s5.r - s4.i;
                    y3.i = s5.i + s4.r;
                }
                cpx fpnk{y3.r, -y3.i};
                cpx f1k = c_add(x0, fpnk), f2k = c_sub(x0, fpnk);
                cpx w = c_mul(f2k, stw[k - 1]);
                float re = (f1k.r + w.r) * 0.5f, im = (f1k.i + w.i) * 0.5f;
                mags[k] = sqrtf(re * re + im * im) * norm_factor;
            }
        }
        __syncthreads();
        // averageVolumeInBand (PipelineFFT.zig:99-112): in-order f32 SUM of bins lo..hi
        if (t < bands.n_bands) {
            float v = 0.0f;
            for (int i = bands.lo[t]; i <= bands.hi[t]; i++) v += mags[i];
            d_band[((size_t)t * n_rows + r) * band_frame_stride + band_frame_offset + f] = v;
        }
        if (FULL && d_spectra != nullptr) {
            float *dst = d_spectra + fid * (size_t)kBins;
            for (int k = t; k < kBins; k += 64) dst[k] = mags[k];
        }
        __syncthreads();
    }
}

int stft_launch(const StftTables &tb, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames,
                size_t hop, const StftBands &bands, float *d_band, size_t band_frame_stride,
                size_t band_frame_offset, float *d_spectra, const float *d_window_override, float norm_override,
                int device, cudaStream_t stream) {
    if (n_rows == 0 || n_frames == 0) return FVAD_OK;
    if ((reinterpret_cast<uintptr_t>(d_pcm) & 7) || (row_stride & 1) || (hop & 1))
        return set_error(FVAD_EINVALID_ARGUMENT, "stft: pcm must be 8-byte aligned with even strides");
    bool full = d_spectra != nullptr;
    StftBands b = bands;
    // the pruned kernel needs every band inside 1..255 and reads the covering range lo[0]..hi[last]
    int lo = 1 << 30, hi = -1;
    for (int i = 0; i < b.n_bands; i++) {
        if (b.lo[i] < 1 || b.hi[i] > 255 || b.hi[i] < b.lo[i]) full = true;
        lo = b.lo[i] < lo ? b.lo[i] : lo;
        hi = b.hi[i] > hi ? b.hi[i] : hi;
    }
    if (b.n_bands == 0) full = true;
    if (!full && hi - lo > 200) full = true;
    const float *win = d_window_override ? d_window_override : tb.d_window;
    const float norm = d_window_override ? norm_override : tb.norm_factor;
    const size_t total = n_rows * n_frames;
    const int sms = device_sm_count(device);
    size_t grid = (size_t)sms * 16;                    // persistent: 16 CTAs of 2 warps per SM (64 registers per thread)
    if (grid > total) grid = total;
    if (full) {
        FVAD_LAUNCH(k_stft2048<true>, (unsigned)grid, 64, 0, stream, d_pcm, n_rows, row_stride, n_frames, hop, win,
                    tb.d_tw, tb.d_stw, tb.d_twB, tb.twA, norm, b, d_band, band_frame_stride, band_frame_offset, d_spectra);
    } else {
        // covering range for the pruned pass: lo[0]..hi[n-1] must span every band
        StftBands bb = b;
        // keep bands as given but make element 0 / last carry the extremes for the kernel's range read
        int imin = 0, imax = 0;
        for (int i = 0; i < b.n_bands; i++) { if (b.lo[i] < b.lo[imin]) imin = i; if (b.hi[i] > b.hi[imax]) imax = i; }
        (void)imin; (void)imax;
        // simplest: pass the range through a dedicated pair by reordering is unnecessary when n_bands==1;
        // for several bands sort so that lo[0] is the smallest lo and hi[last] the largest hi is not
        // generally possible, so fall back to FULL unless the given order already satisfies it.
        bool ok = bb.lo[0] == lo && bb.hi[bb.n_bands - 1] == hi;
        if (!ok) {
            FVAD_LAUNCH(k_stft2048<true>, (unsigned)grid, 64, 0, stream, d_pcm, n_rows, row_stride, n_frames, hop, win,
                        tb.d_tw, tb.d_stw, tb.d_twB, tb.twA, norm, b, d_band, band_frame_stride, band_frame_offset, d_spectra);
        } else {
            FVAD_LAUNCH(k_stft2048<false>, (unsigned)grid, 64, 0, stream, d_pcm, n_rows, row_stride, n_frames, hop,
                        win, tb.d_tw, tb.d_stw, tb.d_twB, tb.twA, norm, bb, d_band, band_frame_stride, band_frame_offset, d_spectra);
        }
    }
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad
