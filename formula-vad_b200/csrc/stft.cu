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
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace fvad {

// ------------------------------------------------------------------------------------------
// host tables
// ------------------------------------------------------------------------------------------
// `@cos` on an f32 in Zig is compiler-rt's cosf, a port of musl's (FreeBSD msun s_cosf.c / k_cosf.c / k_sinf.c): range
// split at odd multiples of pi/4, degree-8 / degree-9 polynomials evaluated in double, one rounding to float.  It is NOT
// correctly rounded: for the 2048-entry periodic Hann window it differs from (float)cos((double)x) in ONE entry (n = 196)
// by one ulp of the cosine.  The window is part of every spectrum, so the table is built with the same function the
// reference calls.  Only |x| <= 9 pi / 4 is needed (the window's arguments lie in [0, 2 pi)).
namespace {
float zig_cosdf(double x) {
    const double C0 = -0x1ffffffd0c5e81.0p-54, C1 = 0x155553e1053a42.0p-57, C2 = -0x16c087e80f1e27.0p-62, C3 = 0x199342e0ee5069.0p-68;
    const double z = x * x, w = z * z, r = C2 + z * C3;
    return (float)(((1.0 + z * C0) + w * C1) + (w * z) * r);
}
float zig_sindf(double x) {
    const double S1 = -0x15555554cbac77.0p-55, S2 = 0x111110896efbb2.0p-59, S3 = -0x1a00f9e2cae774.0p-65, S4 = 0x16cd878c3b46a7.0p-71;
    const double z = x * x, w = z * z, r = S3 + z * S4, s = z * x;
    return (float)((x + s * (S1 + z * S2)) + s * w * r);
}
float zig_cosf(float x) {
    const double pio2 = 1.57079632679489661923;
    uint32_t ix;
    memcpy(&ix, &x, 4);
    const bool sign = (ix >> 31) != 0;
    ix &= 0x7fffffffu;
    if (ix <= 0x3f490fdau) return ix < 0x39800000u ? 1.0f : zig_cosdf(x);                  // |x| <= pi/4
    if (ix <= 0x407b53d1u) {                                                                // |x| <= 5 pi/4
        if (ix > 0x4016cbe3u) return -zig_cosdf(sign ? x + 2 * pio2 : x - 2 * pio2);
        return sign ? zig_sindf(x + pio2) : zig_sindf(pio2 - x);
    }
    if (ix <= 0x40e231d5u) {                                                                // |x| <= 9 pi/4
        if (ix > 0x40afeddfu) return zig_cosdf(sign ? x + 4 * pio2 : x - 4 * pio2);
        return sign ? zig_sindf(-x - 3 * pio2) : zig_sindf(x - 3 * pio2);
    }
    return (float)cos((double)x);                                                           // not reached by the window
}
}  // namespace

// hannWindowPeriodic -> cosineSumWindowPeriodic (window_fn.zig:22-28,51-68): K=1, alphas {0.5,0.5}, evaluated in f32.
void host_hann_window_periodic(float *w, size_t n) {
    const float N = (float)n;
    const float two_pi = (float)(2 * 3.14159265358979323846264338327950288419716939937510);
    for (size_t i = 0; i < n; i++) {
        float nf = (float)i;
        float r = 0;
        r += (1.0f * 0.5f) * zig_cosf(((two_pi * 0.0f) * nf) / N);
        r += (-1.0f * 0.5f) * zig_cosf(((two_pi * 1.0f) * nf) / N);
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
        FVAD_CUDA_TRY(cudaMemcpy(t.d_window, w.data(), sizeof(float) * kFft, cudaMemcpyHostToDevice));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_twB4, sizeof(float2) * 15 * 64));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_tw4, sizeof(float2) * kFftC));
        FVAD_CUDA_TRY(cudaMalloc(&t.d_stw4, sizeof(float2) * kFftC / 2));
        {
            std::vector<cpx> twb(15 * 64);
            for (int th = 0; th < 64; th++) {             // pass-B twiddles per thread t = (pb, jj), [15][64]
                const int jj = th & 15;
                for (int q = 1; q < 4; q++) twb[(q - 1) * 64 + th] = tw[jj * 16 * q];                            // stage m=16
                for (int g = 0; g < 4; g++)
                    for (int q = 1; q < 4; q++) twb[(3 + 3 * g + q - 1) * 64 + th] = tw[(jj + 16 * g) * 4 * q];  // stage m=64
            }
            FVAD_CUDA_TRY(cudaMemcpy(t.d_twB4, twb.data(), sizeof(cpx) * twb.size(), cudaMemcpyHostToDevice));
            FVAD_CUDA_TRY(cudaMemcpy(t.d_tw4, tw.data(), sizeof(cpx) * tw.size(), cudaMemcpyHostToDevice));
            FVAD_CUDA_TRY(cudaMemcpy(t.d_stw4, stw.data(), sizeof(cpx) * stw.size(), cudaMemcpyHostToDevice));
            for (int k = 1; k < 4; k++)
                for (int q = 1; q < 4; q++) {             // stage m=4, uniform over threads: kernel parameter
                    const cpx w = tw[k * 64 * q];
                    t.twA.v[(k - 1) * 3 + q - 1] = make_float2(w.r, w.i);
                }
            t.twA.neg_zero = 0x8000000080000000ull;
        }
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
// One CTA of 64 threads transforms TWO frames at a time with Blackwell's packed f32x2 arithmetic
// (FADD2 / FMUL2; sm_100 __fadd2_rn / __fmul2_rn): every complex value is held as re = (A.re, B.re),
// im = (A.im, B.im) for frames A and B, so both frames execute the same instruction on the same twiddle
// and each IEEE operation is done once per frame exactly as KissFFT does it -- half the issue slots per
// frame, which is what moves this kernel from issue-bound to HBM-bound.
struct c2 { float2 re, im; };                                   // one complex value of frame A and of frame B
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }   // a + (-b) == a - b
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
// ptxas (12.9) contracts a packed multiply feeding a packed add/sub into FFMA2 even for mul.rn.f32x2 and with
// -fmad=false, which would drop a rounding the reference performs.  Products that feed a sum are therefore
// formed as fma(a, b, -0): round(a*b) + (-0) is round(a*b) bit for bit (signed zeros included), it issues in
// the same slot as FMUL2, and an FFMA2 result cannot be contracted into the following FADD2.  The (-0, -0)
// addend is a kernel parameter (StftTwA::neg_zero): given it as a literal, ptxas folds the fma back to a
// multiply and then contracts that.
// tools/check_sass.sh asserts the FFMA2 count of each kernel variant equals the number written here.
__device__ __forceinline__ float2 prod2(float2 a, float2 b, unsigned long long nz) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(*reinterpret_cast<unsigned long long *>(&d))
        : "l"(*reinterpret_cast<const unsigned long long *>(&a)), "l"(*reinterpret_cast<const unsigned long long *>(&b)), "l"(nz));
    return d;
}
__device__ __forceinline__ c2 cadd(c2 a, c2 b) { return c2{add2(a.re, b.re), add2(a.im, b.im)}; }
__device__ __forceinline__ c2 csub(c2 a, c2 b) { return c2{sub2(a.re, b.re), sub2(a.im, b.im)}; }
// C_MUL (kiss_fft _kiss_fft_guts.h): m.r = a.r*b.r - a.i*b.i, m.i = a.r*b.i + a.i*b.r.  The twiddle arrives as
// 8 bytes and is duplicated in registers: the load/store pipe, not the ALU, is this kernel's scarce resource.
__device__ __forceinline__ c2 cmul(c2 a, float2 w, unsigned long long nz) {
    const float2 wr = make_float2(w.x, w.x), wi = make_float2(w.y, w.y);
    return c2{sub2(prod2(a.re, wr, nz), prod2(a.im, wi, nz)), add2(prod2(a.re, wi, nz), prod2(a.im, wr, nz))};
}
// radix-4 butterfly, forward branch of kf_bfly4 (kiss_fft.c:38-84); s0,s1,s2 are the already twiddled
// inputs F[m]*tw1, F[2m]*tw2, F[3m]*tw3.
__device__ __forceinline__ void bfly4(c2 &f0, c2 &f1, c2 &f2, c2 &f3, c2 s0, c2 s1, c2 s2) {
    const c2 s5 = csub(f0, s1);
    f0 = cadd(f0, s1);
    const c2 s3 = cadd(s0, s2);
    const c2 s4 = csub(s0, s2);
    f2 = csub(f0, s3);
    f0 = cadd(f0, s3);
    f1.re = add2(s5.re, s4.im);
    f1.im = sub2(s5.im, s4.re);
    f3.re = sub2(s5.re, s4.im);
    f3.im = add2(s5.im, s4.re);
}

// shared-memory slot of complex position pos (0..1023): one element of padding per 16 and per 256 positions
// keeps all three access patterns (16-blocks, stride-16, stride-256) conflict-free (16-byte elements:
// every quarter-warp touches eight distinct 16-byte bank groups).
__device__ __forceinline__ int slot(int pos) { return pos + (pos >> 4) + (pos >> 8); }
constexpr int kSlots = 1024 + 64 + 4 + 4;
__device__ __forceinline__ c2 lds_c2(const float4 *sm, int pos) { const float4 v = sm[slot(pos)]; return c2{make_float2(v.x, v.y), make_float2(v.z, v.w)}; }
__device__ __forceinline__ void sts_c2(float4 *sm, int pos, c2 v) { sm[slot(pos)] = make_float4(v.re.x, v.re.y, v.im.x, v.im.y); }

// ---- mbarrier + 1-D bulk copy (TMA engine, no tensor map needed for a contiguous 8 KB frame) ----------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// one thread: order the CTA's earlier generic-proxy reads of the staging buffer before the async-proxy refill, arm
// the barrier with the byte count and start both frame copies of pair `pid`
__device__ __forceinline__ void stage_pair(float *dstA, float *dstB, unsigned long long *bar, const float *pcm, size_t rA,
                                           size_t fA, size_t n_rows, size_t n_frames, size_t row_stride, size_t hop) {
    size_t rB = rA, fB = fA + 1;
    if (fB == n_frames) { fB = 0; rB = rA + 1; }
    if (rB >= n_rows) { rB = rA; fB = fA; }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(2u * kFft * 4u) : "memory");
    bulk_g2s(dstA, pcm + rA * row_stride + fA * hop, kFft * 4u, bar);
    bulk_g2s(dstB, pcm + rB * row_stride + fB * hop, kFft * 4u, bar);
}

// LEAN (with FULL): spectra wanted (the bins a band reads are also kept in shared memory for the in-order sums).  The last stage and the kiss_fftr split are fused in registers -- a thread
// takes butterfly kb AND butterfly 256-kb of the last stage, whose eight outputs are exactly the four split pairs
// (k, 1024-k) for k = kb, kb+256, 256-kb, 512-kb -- and the magnitudes go straight from registers to global memory (a
// warp writes 32 consecutive bins per store).  Against the general FULL form that is 16 STS.128 + 18 LDS.128 + 70 4-byte
// shared-memory accesses less per thread and pair (the kernel is bound by the load/store pipe), and no `mags` array, so
// the staged (TMA) input fits beside six resident CTAs.  Arithmetic per output bin is unchanged.
template <bool FULL, bool TMA, bool LEAN = false>
__global__ void __launch_bounds__(64) k_stft2048(const float *__restrict__ pcm, size_t n_rows, size_t row_stride,
                                                 size_t n_frames, size_t hop, const float *__restrict__ window,
                                                 const float2 *__restrict__ tw4, const float2 *__restrict__ stw4,
                                                 const float2 *__restrict__ twB4, const StftTwA twA,
                                                 float norm_factor, StftBands bands, float *__restrict__ d_band,
                                                 size_t band_frame_stride, size_t band_frame_offset,
                                                 float *__restrict__ d_spectra) {
    __shared__ __align__(16) float4 sm[kSlots];
    __shared__ float mags[2][LEAN ? 256 : (FULL ? kBins + 3 : 256)];   // LEAN: only the bins some band reads, [bin - lo[0]]
    // TMA staging: the next pair's two frames (2 x 8 KB, each contiguous in HBM) are fetched by cp.async.bulk into
    // shared memory while this pair is transformed, so no warp ever waits on an HBM load (with 12 warps per SM the
    // register file leaves nothing to hide that latency with).
    __shared__ __align__(128) float stage[TMA ? 2 : 1][TMA ? kFft : 4];
    __shared__ __align__(8) unsigned long long mbar;
    const int t = threadIdx.x;
    const unsigned long long nz = twA.neg_zero;
    // pass B: b = 256-block, jj = position inside the 16-stride comb
    const int pb = t >> 4, jj = t & 15;
    // pass A block: thread t owns input residue n_low = t; its 16 points land in block blk
    const int blk = ((t & 3) << 4) | (t & 12) | (t >> 4);
    const float2 *win2 = reinterpret_cast<const float2 *>(window);
    float2 wreg[16];                                            // this thread's 32 window values stay in registers
#pragma unroll
    for (int e = 0; e < 16; e++) wreg[e] = win2[64 * e + t];
    const size_t total = n_rows * n_frames, pairs = (total + 1) / 2;
    // (row, frame) of this CTA's current frame A, advanced incrementally: a 64-bit division per pair would cost
    // more issue slots than a whole butterfly stage
    const size_t step_r = (2 * (size_t)gridDim.x) / n_frames, step_f = (2 * (size_t)gridDim.x) % n_frames;
    size_t rA = (2 * (size_t)blockIdx.x) / n_frames, fA = (2 * (size_t)blockIdx.x) % n_frames;
    unsigned phase = 0;
    if (TMA) {
        if (t == 0) {
            mbar_init(&mbar, 1);
            if (blockIdx.x < pairs) stage_pair(stage[0], stage[1], &mbar, pcm, rA, fA, n_rows, n_frames, row_stride, hop);
        }
        __syncthreads();
    }

    for (size_t pid = blockIdx.x; pid < pairs; pid += gridDim.x) {
        // frame B = the next frame in (row, frame) order; odd tail: B repeats A and is not stored
        size_t rB = rA, fB = fA + 1;
        if (fB == n_frames) { fB = 0; rB = rA + 1; }
        const bool has_b = rB < n_rows;
        if (!has_b) { rB = rA; fB = fA; }
        size_t rN = rA + step_r, fN = fA + step_f;      // frame A of this CTA's next pair
        if (fN >= n_frames) { fN -= n_frames; rN++; }
        const float2 *srcA, *srcB;
        if (TMA) {
            srcA = reinterpret_cast<const float2 *>(stage[0]);
            srcB = reinterpret_cast<const float2 *>(stage[1]);
            mbar_wait(&mbar, phase);
            phase ^= 1;
        } else {
            srcA = reinterpret_cast<const float2 *>(pcm + rA * row_stride + fA * hop);
            srcB = reinterpret_cast<const float2 *>(pcm + rB * row_stride + fB * hop);
        }

        // ---------------- pass A: load + window, stages m=1 and m=4 in registers ------------
        c2 a[16];
#pragma unroll
        for (int e = 0; e < 16; e++) {
            const float2 vA = srcA[64 * e + t];          // coalesced: 32 lanes x 8 B contiguous
            const float2 vB = srcB[64 * e + t];
            const float2 w = wreg[e];
            // input n = t + 64 e = t + 64 (k4 + 4 k5) -> local slot i = 4 k4 + k5
            const int i = 4 * (e & 3) + (e >> 2);
            a[i].re = make_float2(vA.x * w.x, vB.x * w.x);   // FFT.zig:150,155 sample * window[idx]
            a[i].im = make_float2(vA.y * w.y, vB.y * w.y);
        }
#pragma unroll
        for (int g = 0; g < 4; g++)                     // stage m=1: twiddles are all W^0 (x*(1,-0) == x)
            bfly4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3]);
        bfly4(a[0], a[4], a[8], a[12], a[4], a[8], a[12]);   // stage m=4, k=0
#pragma unroll
        for (int k = 1; k < 4; k++)
            bfly4(a[k], a[k + 4], a[k + 8], a[k + 12], cmul(a[k + 4], twA.v[(k - 1) * 3], nz), cmul(a[k + 8], twA.v[(k - 1) * 3 + 1], nz),
                  cmul(a[k + 12], twA.v[(k - 1) * 3 + 2], nz));
#pragma unroll
        for (int i = 0; i < 16; i++) sts_c2(sm, 16 * blk + i, a[i]);
        __syncthreads();
        if (TMA && t == 0 && pid + gridDim.x < pairs)   // every thread has consumed the staged frames: refill
            stage_pair(stage[0], stage[1], &mbar, pcm, rN, fN, n_rows, n_frames, row_stride, hop);

        // ---------------- pass B: stages m=16 and m=64 -------------------------------------
        const float2 *twb = twB4 + t;
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = lds_c2(sm, 256 * pb + jj + 16 * i);
        {
            const float2 w1 = __ldg(twb), w2 = __ldg(twb + 64), w3 = __ldg(twb + 128);
#pragma unroll
            for (int g = 0; g < 4; g++)                 // m=16: points jj+16*(4g+r), twiddle index jj
                bfly4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3], cmul(a[4 * g + 1], w1, nz), cmul(a[4 * g + 2], w2, nz),
                      cmul(a[4 * g + 3], w3, nz));
        }
#pragma unroll
        for (int g = 0; g < 4; g++) {                   // m=64: points (jj+16g)+64r, twiddle index jj+16g
            const float2 w1 = __ldg(twb + (3 + 3 * g) * 64), w2 = __ldg(twb + (4 + 3 * g) * 64), w3 = __ldg(twb + (5 + 3 * g) * 64);
            bfly4(a[g], a[g + 4], a[g + 8], a[g + 12], cmul(a[g + 4], w1, nz), cmul(a[g + 8], w2, nz), cmul(a[g + 12], w3, nz));
        }
#pragma unroll
        for (int i = 0; i < 16; i++) sts_c2(sm, 256 * pb + jj + 16 * i, a[i]);
        __syncthreads();

        // ---------------- pass C: stage m=256 + kiss_fftr split + |X|*norm -----------------
        const float2 half2 = make_float2(0.5f, 0.5f);
        if (LEAN) {
            float *dstA = d_spectra + (rA * n_frames + fA) * (size_t)kBins;
            float *dstB = d_spectra + (rB * n_frames + fB) * (size_t)kBins;
            const int blo = bands.n_bands ? bands.lo[0] : 0, bspan = bands.n_bands ? bands.hi[bands.n_bands - 1] - blo : -1;
            auto put = [&](int bin, float2 m2) {                     // |X| * norm (FFT.zig:167-176) of frames A and B
                const float mA = sqrtf(m2.x) * norm_factor, mB = sqrtf(m2.y) * norm_factor;
                dstA[bin] = mA;
                if (has_b) dstB[bin] = mB;
                if ((unsigned)(bin - blo) <= (unsigned)bspan) { mags[0][bin - blo] = mA; mags[1][bin - blo] = mB; }   // for the band sums
            };
            auto last_stage = [&](int kb, c2 &f0, c2 &f1, c2 &f2, c2 &f3) {   // tmp[kb + 256 j], j = 0..3 (kiss_fft.c:38-84, m = 256)
                f0 = lds_c2(sm, kb); f1 = lds_c2(sm, kb + 256); f2 = lds_c2(sm, kb + 512); f3 = lds_c2(sm, kb + 768);
                bfly4(f0, f1, f2, f3, cmul(f1, __ldg(tw4 + kb), nz), cmul(f2, __ldg(tw4 + 2 * kb), nz), cmul(f3, __ldg(tw4 + 3 * kb), nz));
            };
            auto split = [&](int k, const c2 &pk, const c2 &qk, bool low) {   // kiss_fftr.c:90-116: p = tmp[k], q = tmp[1024 - k]
                const c2 f1k{add2(pk.re, qk.re), sub2(pk.im, qk.im)}, f2k{sub2(pk.re, qk.re), add2(pk.im, qk.im)};
                const c2 w = cmul(f2k, __ldg(stw4 + (k - 1)), nz);
                const float2 re2 = mul2(sub2(f1k.re, w.re), half2), im2 = mul2(sub2(w.im, f1k.im), half2);
                put(1024 - k, add2(prod2(re2, re2, nz), prod2(im2, im2, nz)));
                if (low) {                                               // k == 512: freqdata[512] is written twice, the second wins
                    const float2 re = mul2(add2(f1k.re, w.re), half2), im = mul2(add2(f1k.im, w.im), half2);
                    put(k, add2(prod2(re, re, nz), prod2(im, im, nz)));
                }
            };
#pragma unroll 1
            for (int it = 0; it < 2; it++) {
                const int kb = t + 1 + 64 * it;                          // 1..127; thread 63's second turn takes 0 and 128
                c2 a0, a1, a2, a3, b0, b1, b2, b3;
                if (kb < 128) {
                    last_stage(kb, a0, a1, a2, a3);
                    last_stage(256 - kb, b0, b1, b2, b3);
                    split(kb, a0, b3, true);                             // partner 1024-kb = (256-kb) + 768
                    split(kb + 256, a1, b2, true);                       // partner 768-kb
                    split(256 - kb, b0, a3, true);                       // partner 768+kb
                    split(512 - kb, b1, a2, true);                       // partner 512+kb
                } else {
                    last_stage(0, a0, a1, a2, a3);
                    last_stage(128, b0, b1, b2, b3);
                    const float2 r0 = add2(a0.re, a0.im), rn = sub2(a0.re, a0.im);     // DC / Nyquist, kiss_fftr.c:84-88
                    const float2 z = make_float2(0.0f * 0.0f, 0.0f * 0.0f);
                    put(0, add2(prod2(r0, r0, nz), z));
                    put(1024, add2(prod2(rn, rn, nz), z));
                    split(256, a1, a3, true);
                    split(512, a2, a2, false);
                    split(128, b0, b3, true);
                    split(384, b1, b2, true);
                }
            }
        } else if (FULL) {
#pragma unroll
            for (int it = 0; it < 4; it++) {
                const int k = t + 64 * it;
                c2 f0 = lds_c2(sm, k), f1 = lds_c2(sm, k + 256), f2 = lds_c2(sm, k + 512), f3 = lds_c2(sm, k + 768);
                bfly4(f0, f1, f2, f3, cmul(f1, __ldg(tw4 + k), nz), cmul(f2, __ldg(tw4 + 2 * k), nz), cmul(f3, __ldg(tw4 + 3 * k), nz));
                sts_c2(sm, k, f0); sts_c2(sm, k + 256, f1); sts_c2(sm, k + 512, f2); sts_c2(sm, k + 768, f3);
            }
            __syncthreads();
            for (int k = t; k <= 512; k += 64) {        // kiss_fftr.c:90-116
                if (k == 0) {
                    const c2 d = lds_c2(sm, 0);
                    const float2 r0 = add2(d.re, d.im), rn = sub2(d.re, d.im);
                    const float2 z = make_float2(0.0f * 0.0f, 0.0f * 0.0f);
                    const float2 q0 = add2(prod2(r0, r0, nz), z), qn = add2(prod2(rn, rn, nz), z);
                    mags[0][0] = sqrtf(q0.x) * norm_factor; mags[1][0] = sqrtf(q0.y) * norm_factor;
                    mags[0][1024] = sqrtf(qn.x) * norm_factor; mags[1][1024] = sqrtf(qn.y) * norm_factor;
                } else {
                    const c2 p = lds_c2(sm, k), q = lds_c2(sm, 1024 - k);
                    // fpnk = conj(q): f1k = p + fpnk, f2k = p - fpnk
                    const c2 f1k{add2(p.re, q.re), sub2(p.im, q.im)}, f2k{sub2(p.re, q.re), add2(p.im, q.im)};
                    const c2 w = cmul(f2k, __ldg(stw4 + (k - 1)), nz);
                    const float2 re = mul2(add2(f1k.re, w.re), half2), im = mul2(add2(f1k.im, w.im), half2);
                    const float2 re2 = mul2(sub2(f1k.re, w.re), half2), im2 = mul2(sub2(w.im, f1k.im), half2);
                    // k == 512: the reference writes freqdata[512] twice, the second wins
                    const float2 m2 = add2(prod2(re2, re2, nz), prod2(im2, im2, nz));
                    mags[0][1024 - k] = sqrtf(m2.x) * norm_factor; mags[1][1024 - k] = sqrtf(m2.y) * norm_factor;
                    if (k != 512) {
                        const float2 m1 = add2(prod2(re, re, nz), prod2(im, im, nz));      // FFT.zig:167-176
                        mags[0][k] = sqrtf(m1.x) * norm_factor; mags[1][k] = sqrtf(m1.y) * norm_factor;
                    }
                }
            }
        } else {
            // only the bins some band reads (all within 1..255)
            const int lo = bands.lo[0], hi = bands.hi[bands.n_bands - 1];   // host guarantees a covering range
            for (int k = lo + t; k <= hi; k += 64) {
                c2 x0, y3;
                {   // output 0 of butterfly k = tmp[k]
                    const c2 v0 = lds_c2(sm, k);
                    const c2 s0 = cmul(lds_c2(sm, k + 256), __ldg(tw4 + k), nz), s1 = cmul(lds_c2(sm, k + 512), __ldg(tw4 + 2 * k), nz),
                             s2 = cmul(lds_c2(sm, k + 768), __ldg(tw4 + 3 * k), nz);
                    x0 = cadd(cadd(v0, s1), cadd(s0, s2));
                }
                {   // output 3 of butterfly 256-k = tmp[1024-k]
                    const int kk = 256 - k;
                    const c2 v0 = lds_c2(sm, kk);
                    const c2 s0 = cmul(lds_c2(sm, kk + 256), __ldg(tw4 + kk), nz), s1 = cmul(lds_c2(sm, kk + 512), __ldg(tw4 + 2 * kk), nz),
                             s2 = cmul(lds_c2(sm, kk + 768), __ldg(tw4 + 3 * kk), nz);
                    const c2 s5 = csub(v0, s1), s4 = csub(s0, s2);
                    y3.re = sub2(s5.re, s4.im);
                    y3.im = add2(s5.im, s4.re);
                }
                const c2 f1k{add2(x0.re, y3.re), sub2(x0.im, y3.im)}, f2k{sub2(x0.re, y3.re), add2(x0.im, y3.im)};
                const c2 w = cmul(f2k, __ldg(stw4 + (k - 1)), nz);
                const float2 re = mul2(add2(f1k.re, w.re), half2), im = mul2(add2(f1k.im, w.im), half2);
                const float2 m1 = add2(prod2(re, re, nz), prod2(im, im, nz));
                mags[0][k] = sqrtf(m1.x) * norm_factor; mags[1][k] = sqrtf(m1.y) * norm_factor;
            }
        }
        __syncthreads();
        // averageVolumeInBand (PipelineFFT.zig:99-112): in-order f32 SUM of bins lo..hi, for frame A and frame B
        if (LEAN && t < 2 * bands.n_bands) {
            const int fr = t >= bands.n_bands, bd = t - fr * bands.n_bands, blo = bands.lo[0];
            if (!fr || has_b) {
                float v = 0.0f;
                for (int i = bands.lo[bd]; i <= bands.hi[bd]; i++) v += mags[fr][i - blo];
                d_band[((size_t)bd * n_rows + (fr ? rB : rA)) * band_frame_stride + band_frame_offset + (fr ? fB : fA)] = v;
            }
        }
        if (!LEAN && t < 2 * bands.n_bands) {
            const int fr = t >= bands.n_bands, bd = t - fr * bands.n_bands;
            if (!fr || has_b) {
                float v = 0.0f;
                for (int i = bands.lo[bd]; i <= bands.hi[bd]; i++) v += mags[fr][i];
                d_band[((size_t)bd * n_rows + (fr ? rB : rA)) * band_frame_stride + band_frame_offset + (fr ? fB : fA)] = v;
            }
        }
        if (FULL && !LEAN && d_spectra != nullptr) {
            float *dstA = d_spectra + (rA * n_frames + fA) * (size_t)kBins;
            for (int k = t; k < kBins; k += 64) dstA[k] = mags[0][k];
            if (has_b) {
                float *dstB = d_spectra + (rB * n_frames + fB) * (size_t)kBins;
                for (int k = t; k < kBins; k += 64) dstB[k] = mags[1][k];
            }
        }
        __syncthreads();
        rA = rN; fA = fN;
    }
}

namespace {
struct StftGrid { int occ[2][2] = {{0, 0}, {0, 0}}; int occ_lean[2] = {0, 0}; int sms = 0; bool ready = false; };
StftGrid g_grid[64];

template <bool FULL, bool TMA, bool LEAN = false>
int stft_occupancy(int *occ) {
    // the staged variant needs the whole shared-memory carve-out to keep six CTAs (12 warps) per SM
    FVAD_CUDA_TRY(cudaFuncSetAttribute(k_stft2048<FULL, TMA, LEAN>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    FVAD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k_stft2048<FULL, TMA, LEAN>, 64, 0));
    if (*occ < 1) *occ = 1;
    return FVAD_OK;
}
}  // namespace

int stft_launch(const StftTables &tb, const float *d_pcm, size_t n_rows, size_t row_stride, size_t n_frames,
                size_t hop, const StftBands &bands, float *d_band, size_t band_frame_stride,
                size_t band_frame_offset, float *d_spectra, const float *d_window_override, float norm_override,
                int device, cudaStream_t stream) {
    if (n_rows == 0 || n_frames == 0) return FVAD_OK;
    if ((reinterpret_cast<uintptr_t>(d_pcm) & 7) || (row_stride & 1) || (hop & 1))
        return set_error(FVAD_EINVALID_ARGUMENT, "stft: pcm must be 8-byte aligned with even strides");
    // cp.async.bulk wants 16-byte aligned sources; anything else takes the plain-load variant
    bool tma = !(reinterpret_cast<uintptr_t>(d_pcm) & 15) && !(row_stride & 3) && !(hop & 3);
    if (const char *e = debug_knob("stft_tma")) tma = tma && atoi(e) != 0;
    // The pruned variant computes only the bins lo[0]..hi[last]; it needs every band inside 1..255 and the first /
    // last band to carry the extremes (true for the pipeline's single speech band).  Otherwise: all 1025 bins.
    bool full = d_spectra != nullptr || bands.n_bands == 0;
    for (int i = 0; i < bands.n_bands && !full; i++)
        full = bands.lo[i] < 1 || bands.hi[i] > 255 || bands.hi[i] < bands.lo[i] || bands.lo[i] < bands.lo[0] ||
               bands.hi[i] > bands.hi[bands.n_bands - 1];
    if (!full && bands.hi[bands.n_bands - 1] - bands.lo[0] > 200) full = true;
    const float *win = d_window_override ? d_window_override : tb.d_window;
    const float norm = d_window_override ? norm_override : tb.norm_factor;
    const size_t total = n_rows * n_frames, pairs = (total + 1) / 2;
    if (device < 0 || device >= 64) return set_error(FVAD_EINVALID_ARGUMENT, "stft: bad device %d", device);
    StftGrid &g = g_grid[device];
    {
        std::lock_guard<std::mutex> lock(g_tab_mu);
        if (!g.ready) {
            FVAD_CUDA_TRY(cudaDeviceGetAttribute(&g.sms, cudaDevAttrMultiProcessorCount, device));
            int rc;
            if ((rc = stft_occupancy<false, false>(&g.occ[0][0])) != FVAD_OK) return rc;
            if ((rc = stft_occupancy<false, true>(&g.occ[0][1])) != FVAD_OK) return rc;
            if ((rc = stft_occupancy<true, false>(&g.occ[1][0])) != FVAD_OK) return rc;
            if ((rc = stft_occupancy<true, false, true>(&g.occ_lean[0])) != FVAD_OK) return rc;
            if ((rc = stft_occupancy<true, true, true>(&g.occ_lean[1])) != FVAD_OK) return rc;
            g.ready = true;
        }
    }
    // spectra wanted (FFT.fft, Spectrogram.compute, fvad_stft_batch(d_spectra)) and the bands, if any, ordered and within 256 bins
    bool lean = full && d_spectra != nullptr;
    for (int i = 0; i < bands.n_bands && lean; i++)
        lean = bands.hi[i] >= bands.lo[i] && bands.lo[i] >= bands.lo[0] && bands.hi[i] <= bands.hi[bands.n_bands - 1] &&
               bands.hi[bands.n_bands - 1] - bands.lo[0] < 256;
    if (full && !lean) tma = false;           // the general form's `mags` array plus the staging buffer would cost a resident CTA
    size_t grid = (size_t)g.sms * (lean ? g.occ_lean[tma] : g.occ[full][tma]);
    if (grid > pairs) grid = pairs;
#define FVAD_STFT_GO(F, T)                                                                                              \
    FVAD_LAUNCH_NAMED("k_stft2048<" #F ">", (k_stft2048<F, T>), (unsigned)grid, 64, 0, stream, d_pcm, n_rows, row_stride, n_frames, hop, win, tb.d_tw4, \
                tb.d_stw4, tb.d_twB4, tb.twA, norm, bands, d_band, band_frame_stride, band_frame_offset, d_spectra)
#define FVAD_STFT_LEAN(T)                                                                                               \
    FVAD_LAUNCH_NAMED("k_stft2048<true>", (k_stft2048<true, T, true>), (unsigned)grid, 64, 0, stream, d_pcm, n_rows, row_stride, n_frames, hop, win, tb.d_tw4, \
                tb.d_stw4, tb.d_twB4, tb.twA, norm, bands, d_band, band_frame_stride, band_frame_offset, d_spectra)
    if (lean)      { if (tma) FVAD_STFT_LEAN(true); else FVAD_STFT_LEAN(false); }
    else if (full) FVAD_STFT_GO(true, false);
    else           { if (tma) FVAD_STFT_GO(false, true); else FVAD_STFT_GO(false, false); }
#undef FVAD_STFT_LEAN
#undef FVAD_STFT_GO
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad
