// What bounds the biquad chain lane?  One warp, registers only, clock64().  (round 2)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false -o chain_mix chain_mix.cu && ./chain_mix
#include "../../formula-vad_b200/csrc/rn_core.cuh"
#include <cstdio>
#include <cstdlib>
using namespace fvad::rn;
constexpr int ITER = 4096, U = 16;

__device__ __forceinline__ void step6(float x, float &m0, float &m1) {      // w by one FFMA (NOT robust under cancellation): 6-op chain
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float dh = fma_rn(-2.0f, x, m1);
    const float bb = dh - m1;
    const float dl = (m1 - (dh - bb)) + fma_rn(-2.0f, x, -bb);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    const float w = fma_rn(A0, y, dh);
    const bool pbig = fabsf(y) >= fabsf(dh) * (1.0f / 1.99599f);
    const float hi = pbig ? p : dh, lo = pbig ? dh : p;
    const float f = (hi - w) + lo;
    const float uh = (pl + dl) + f;
    m1 = fma_rn(-A1, y, x);
    m0 = w + uh;
}
__device__ __forceinline__ void step_t(float x, float &m0, float &m1) {     // cancellation form: d = (p + dh) - w', no compare / selects
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float dh = fma_rn(-2.0f, x, m1);
    const float bb = dh - m1;
    const float dl = (m1 - (dh - bb)) + fma_rn(-2.0f, x, -bb);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    const float w = fma_rn(A0, y, dh);
    const float t = p + dh;
    const float u = (t - w) + (pl + dl);
    m1 = fma_rn(-A1, y, x);
    m0 = w + u;
}
__device__ __forceinline__ void step_nodl(float x, float &m0, float &m1) {  // v2 without the TwoSum of (m1, -2x): how much is its issue cost?
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float dh = fma_rn(-2.0f, x, m1);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    const float w = p + dh;
    const bool pbig = fabsf(y) >= fabsf(dh) * (1.0f / 1.99599f);
    const float hi = pbig ? p : dh, lo = pbig ? dh : p;
    const float f = (hi - w) + lo;
    const float uh = pl + f;
    m1 = fma_rn(-A1, y, x);
    m0 = w + uh;
}
__device__ __forceinline__ void step_chain7(float x, float &m0, float &m1) { // ONLY the 7 dependent operations
    const float y = x + m0;
    const float p = 1.99599f * y;
    const float w = p + m1;
    const float f1 = x - w;
    const float f = f1 + m1;
    const float uh = x + f;
    m0 = w + uh;
}
__device__ __forceinline__ void step_chain4(float x, float &m0, float &m1) { // 4 dependent operations
    const float y = x + m0;
    const float p = 1.99599f * y;
    const float w = p + m1;
    m0 = w + x;
}

__device__ __forceinline__ void step_pred(float x, float &m0, float &m1) {  // v2 with the Fast2Sum operands chosen by PREDICATED adds, no FSEL
    const float A0 = 1.99599f, A1 = 0.99600f;
    const float dh = fma_rn(-2.0f, x, m1);
    const float bb = dh - m1;
    const float dl = (m1 - (dh - bb)) + fma_rn(-2.0f, x, -bb);
    const float y = x + m0;
    const float p = A0 * y, pl = fma_rn(A0, y, -p);
    const float w = p + dh;
    const float k = fabsf(dh) * (1.0f / 1.99599f);
    float f;
    asm("{\n\t.reg .pred q;\n\t.reg .f32 t;\n\t"
        "setp.ge.f32 q, %1, %2;\n\t"
        "@q sub.rn.f32 t, %3, %5;\n\t"
        "@!q sub.rn.f32 t, %4, %5;\n\t"
        "@q add.rn.f32 %0, t, %4;\n\t"
        "@!q add.rn.f32 %0, t, %3;\n\t}"
        : "=f"(f) : "f"(fabsf(y)), "f"(k), "f"(p), "f"(dh), "f"(w));
    const float uh = (pl + dl) + f;
    m1 = fma_rn(-A1, y, x);
    m0 = w + uh;
}

// memory variants: the chain lane with the kernel's shared-memory traffic (inputs prefetched by LDS.128 one iteration ahead,
// state snapshots stored by STS.128), one warp
template <int V> __global__ void __launch_bounds__(32) runm(const float *__restrict__ x, float *__restrict__ out, long long *cyc) {
    __shared__ __align__(16) float sx[1024];
    __shared__ __align__(16) float sp[2048];
    __shared__ __align__(16) float sy[1024];
    for (int i = threadIdx.x; i < 1024; i += 32) sx[i] = x[i & 31];
    __syncwarp();
    float m0 = x[40], m1 = x[41];
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
#pragma unroll 1
        for (int blk = 0; blk < ITER * U / 96; blk++) {
            const float *xin = &sx[(blk & 7) * 96];
            float4 xc[4], xn[4];
#pragma unroll
            for (int k = 0; k < 4; k++) xc[k] = *reinterpret_cast<const float4 *>(xin + 4 * k);
#pragma unroll 1
            for (int i = 0; i < 96; i += 16) {
                const int inx = i + 16 < 96 ? i + 16 : i;
#pragma unroll
                for (int k = 0; k < 4; k++) xn[k] = *reinterpret_cast<const float4 *>(xin + inx + 4 * k);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    if (V == 0 || V == 1 || V == 4) {              // v2 step; V1: + one STS.128 of m0 snapshots; V4: predicated form + STS
                        float4 pm;
                        if (V == 4) {
                            pm.x = m0; step_pred(xc[k].x, m0, m1); pm.y = m0; step_pred(xc[k].y, m0, m1);
                            pm.z = m0; step_pred(xc[k].z, m0, m1); pm.w = m0; step_pred(xc[k].w, m0, m1);
                        } else {
                            pm.x = m0; biquad_chain_step(xc[k].x, m0, m1); pm.y = m0; biquad_chain_step(xc[k].y, m0, m1);
                            pm.z = m0; biquad_chain_step(xc[k].z, m0, m1); pm.w = m0; biquad_chain_step(xc[k].w, m0, m1);
                        }
                        if (V >= 1) *reinterpret_cast<float4 *>(&sp[i + 4 * k]) = pm;
                    } else {                                       // round-1 form: three STS.128 per four samples
                        float4 sa, sb, yy;
                        const float sc = 1.0f / 32767.0f;
                        sa.x = m0; sa.y = m1; yy.x = biquad_step_fast(xc[k].x * sc, m0, m1);
                        sa.z = m0; sa.w = m1; yy.y = biquad_step_fast(xc[k].y * sc, m0, m1);
                        sb.x = m0; sb.y = m1; yy.z = biquad_step_fast(xc[k].z * sc, m0, m1);
                        sb.z = m0; sb.w = m1; yy.w = biquad_step_fast(xc[k].w * sc, m0, m1);
                        if (V == 3) {
                            *reinterpret_cast<float4 *>(&sp[2 * (i + 4 * k)]) = sa;
                            *reinterpret_cast<float4 *>(&sp[2 * (i + 4 * k) + 4]) = sb;
                            *reinterpret_cast<float4 *>(&sy[i + 4 * k]) = yy;
                        }
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; k++) xc[k] = xn[k];
            }
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = m0 + m1 + sp[threadIdx.x] + sy[threadIdx.x];
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// block structure of k_rn_biquad_v2 (96 samples per block, 128 threads): where do 33 -> 42 cycles per sample come from?
//   V=0: one warp, blocks of 96 with prologue / epilogue, no barriers      V=1: + 3 idle warps and the two barriers per block
//   V=2: + the verifying threads' work                                     V=3: V=2 with 192-sample blocks
template <int V, int BQ> __global__ void __launch_bounds__(32 + ((BQ + 31) / 32) * 32) runb(const float *__restrict__ x, float *__restrict__ out, long long *cyc, int nblk) {
    __shared__ __align__(16) float sx[2][960];
    __shared__ __align__(16) float sy[2][960];
    __shared__ __align__(16) float PM[2][BQ + 4];
    __shared__ float M1[2][2];
    __shared__ int FINAL[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 1920; i += blockDim.x) (&sx[0][0])[i] = x[i & 31];
    if (threadIdx.x < 2) FINAL[threadIdx.x] = 0;
    __syncthreads();
    float m0 = x[40], m1 = x[41];
    const long long t0 = clock64();
    for (int g = 0; g <= nblk + 1; g++) {
        bool hz = false;
        if (warp == 0) {
            if (lane == 0 && g < nblk) {
                const int par = g & 1, b0 = (g % (960 / BQ)) * BQ;
                const float *xin = &sx[par][b0];
                const float a0 = fabsf(m0), a1 = fabsf(m1);
                if ((a0 != 0.0f && a0 < 1.0e-25f) || (a1 != 0.0f && a1 < 1.0e-25f)) {
                    for (int i = 0; i < BQ; i++) sy[par][b0 + i] = biquad_step_x(xin[i], m0, m1);
                    FINAL[par] = 1;
                } else {
                    M1[par][0] = m1;
                    float4 xc[4], xn[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) xc[k] = *reinterpret_cast<const float4 *>(xin + 4 * k);
#pragma unroll 1
                    for (int i = 0; i < BQ; i += 16) {
                        const int inx = i + 16 < BQ ? i + 16 : i;
#pragma unroll
                        for (int k = 0; k < 4; k++) xn[k] = *reinterpret_cast<const float4 *>(xin + inx + 4 * k);
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            float4 pm;
                            pm.x = m0; biquad_chain_step(xc[k].x, m0, m1); pm.y = m0; biquad_chain_step(xc[k].y, m0, m1);
                            pm.z = m0; biquad_chain_step(xc[k].z, m0, m1); pm.w = m0; biquad_chain_step(xc[k].w, m0, m1);
                            *reinterpret_cast<float4 *>(&PM[par][i + 4 * k]) = pm;
                        }
#pragma unroll
                        for (int k = 0; k < 4; k++) xc[k] = xn[k];
                    }
                    PM[par][BQ] = m0; M1[par][1] = m1; FINAL[par] = 0;
                }
            }
        } else if (V >= 2) {
            const int v = (int)threadIdx.x - 32;
            if (g >= 1 && g - 1 < nblk && v < BQ) {
                const int gv = g - 1, par = gv & 1, b0 = (gv % (960 / BQ)) * BQ;
                if (!FINAL[par]) {
                    const float *xb = &sx[par][b0];
                    float a = PM[par][v], b;
                    if (v == 0) b = M1[par][0];
                    else { const float xp = xb[v - 1]; b = __fmaf_rn(-0.99600f, xp + PM[par][v - 1], xp); }
                    const float xv = xb[v];
                    const float yv = biquad_step_eft_x(xv, a, b, hz);
                    hz |= (__float_as_uint(a) != __float_as_uint(PM[par][v + 1])) | (__float_as_uint(b) != __float_as_uint(__fmaf_rn(-0.99600f, yv, xv)));
                    sy[par][b0 + v] = yv;
                }
            }
        }
        if (V >= 1) {
            const int verdict = __syncthreads_or(hz ? 1 : 0);
            if (threadIdx.x == 0 && verdict && g < 0) m0 += 1.0f;      // (no replay in the probe)
            __syncthreads();
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = m0 + m1 + sy[0][threadIdx.x];
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int V> __global__ void __launch_bounds__(32) run(const float *__restrict__ x, float *__restrict__ out, long long *cyc) {
    float xr[U];
    for (int i = 0; i < U; i++) xr[i] = x[i];
    float m0 = x[40], m1 = x[41];
    float acc[8];
    for (int i = 0; i < 8; i++) acc[i] = x[48 + i];
    const float c0 = x[60], c1 = x[61];
    __syncwarp();
    const long long t0 = clock64();
    if (threadIdx.x == 0 || V >= 20) {
#pragma unroll 1
        for (int it = 0; it < ITER; it++) {
#pragma unroll
            for (int i = 0; i < U; i++) {
                if (V == 0) biquad_step_fast(xr[i] * (1.0f / 32767.0f), m0, m1);
                if (V == 1) biquad_chain_step(xr[i], m0, m1);
                if (V == 2) step6(xr[i], m0, m1);
                if (V == 3) step_t(xr[i], m0, m1);
                if (V == 4) step_nodl(xr[i], m0, m1);
                if (V == 5) step_chain7(xr[i], m0, m1);
                if (V == 6) step_chain4(xr[i], m0, m1);
                if (V == 13) step_pred(xr[i], m0, m1);
                if (V == 7 || V == 20) {                       // 8 independent FADD R,R,R per "sample"
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] = acc[k] + xr[i];
                }
                if (V == 8) {                                  // 8 independent FFMA R,R,R,R
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] = fma_rn(acc[k], c0, xr[i]);
                }
                if (V == 9) {                                  // 8 independent FFMA with an immediate
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] = fma_rn(acc[k], 0.996f, xr[i]);
                }
                if (V == 10) {                                 // 8 independent FMUL with an immediate
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] = acc[k] * 0.999f;
                }
                if (V == 11) {                                 // 4 FADD + 4 FFMA-imm, independent
#pragma unroll
                    for (int k = 0; k < 4; k++) { acc[k] = acc[k] + xr[i]; acc[k + 4] = fma_rn(acc[k + 4], 0.996f, c1); }
                }
                if (V == 12) {                                 // 8 independent FSEL on a fixed predicate + FADD feeding
#pragma unroll
                    for (int k = 0; k < 8; k++) acc[k] = acc[k] > c0 ? acc[k] : xr[i];
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = m0 + m1;
    for (int i = 0; i < 8; i++) s += acc[i];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

int main() {
    float hx[64];
    unsigned s = 12345;
    for (int i = 0; i < 64; i++) { s = s * 1664525u + 1013904223u; hx[i] = ((int)(s >> 8) - (1 << 23)) / (float)(1 << 24) * 0.06f * 32767.0f; }
    hx[40] = 0.01f; hx[41] = -0.02f;
    float *x, *out; long long *cyc, c;
    cudaMalloc(&x, sizeof hx); cudaMalloc(&out, 128); cudaMalloc(&cyc, 8);
    cudaMemcpy(x, hx, sizeof hx, cudaMemcpyHostToDevice);
    const double n = (double)ITER * U;
#define R(V, name, per) for (int rep = 0; rep < 2; rep++) { run<V><<<1, 32>>>(x, out, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); \
        if (rep) printf("%-58s: %6.2f cyc/%s\n", name, (double)c / n / (per), (per) == 1 ? "sample" : "instr"); }
    R(0, "biquad_step_fast (round 1 chain step)", 1)
    R(1, "biquad_chain_step (v2: scaled input, folded -2x)", 1)
    R(2, "w by FFMA, 6-op chain (not robust)", 1)
    R(3, "cancellation form, no compare/selects (not robust)", 1)
    R(4, "v2 without TwoSum(m1,-2x) (wrong, issue probe)", 1)
    R(5, "only 7 dependent ops", 1)
    R(6, "only 4 dependent ops", 1)
    R(7, "8 independent FADD R,R,R, lane 0 only", 8)
    R(20, "8 independent FADD R,R,R, whole warp", 8)
    R(8, "8 independent FFMA R,R,R,R", 8)
    R(9, "8 independent FFMA imm", 8)
    R(10, "8 independent FMUL imm", 8)
    R(11, "4 FADD + 4 FFMA imm", 8)
    R(12, "8 independent FSETP+FSEL", 8)
    R(13, "v2 with predicated adds instead of FSEL", 1)
    const double nm = (double)(ITER * U / 96) * 96;
#define RM(V, name) for (int rep = 0; rep < 2; rep++) { runm<V><<<1, 32>>>(x, out, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); \
        if (rep) printf("%-58s: %6.2f cyc/sample\n", name, (double)c / nm); }
    RM(0, "mem: v2 step, LDS.128 prefetch, no stores")
    RM(1, "mem: v2 step, LDS.128 prefetch, 1 STS.128 / 4 samples")
    RM(2, "mem: round-1 step, LDS.128 prefetch, no stores")
    RM(3, "mem: round-1 step, LDS.128 prefetch, 3 STS.128 / 4 samples")
    RM(4, "mem: predicated step, LDS.128 prefetch, 1 STS.128 / 4")
    const int nblk = 6000;
#define RB(V, BQ, T, name) for (int rep = 0; rep < 2; rep++) { runb<V, BQ><<<1, T>>>(x, out, cyc, nblk); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); \
        if (rep) printf("%-58s: %6.2f cyc/sample\n", name, (double)c / ((double)nblk * BQ)); }
    RB(0, 96, 32, "blocks of 96: one warp, no barriers")
    RB(1, 96, 128, "blocks of 96: 4 warps, two barriers per block")
    RB(2, 96, 128, "blocks of 96: + verification work")
    RB(2, 192, 224, "blocks of 192: + verification work")
    RB(2, 48, 96, "blocks of 48: + verification work")
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
