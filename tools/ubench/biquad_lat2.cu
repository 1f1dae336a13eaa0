// Where do the cycles of the f32 (error-free transformation) biquad go?  One warp, clock64().
#include "../../formula-vad_b200/csrc/rn_core.cuh"
#include <cstdio>
#include <cstdlib>
using namespace fvad::rn;
constexpr int NB = 1500;   // blocks of 32
template <int V> __global__ void __launch_bounds__(32) run(const float *__restrict__ x, float *__restrict__ y, long long *cyc) {
    __shared__ float sx[32], sy[32], sm0[33], sm1[33];
    __shared__ int flag;
    const int lane = threadIdx.x;
    if (lane == 0) flag = 0;
    __syncwarp();
    long long t0 = clock64();
    float m0 = 0, m1 = 0;
    for (int b = 0; b < NB; b++) {
        sx[lane] = x[b * 32 + lane];
        __syncwarp();
        if (V == 0) {                       // chain only, registers
            if (lane == 0) {
                float xr[32];
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = sx[i];
                float acc = 0;
#pragma unroll
                for (int i = 0; i < 32; i++) acc += biquad_step_fast(xr[i], m0, m1);
                sy[0] = acc;
            }
        } else if (V == 1) {                // chain + y stores
            if (lane == 0) {
                float xr[32];
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = sx[i];
#pragma unroll
                for (int i = 0; i < 32; i++) sy[i] = biquad_step_fast(xr[i], m0, m1);
            }
        } else if (V == 2) {                // full verified block
            biquad_block_warp(sx, sy, m0, m1, sm0, sm1, &flag);
        } else if (V == 4 || V == 5 || V == 6) {   // chain only, smaller unrolled bodies (I-cache hypothesis)
            constexpr int U = V == 4 ? 4 : V == 5 ? 8 : 16;
            if (lane == 0) {
                float acc = 0;
#pragma unroll 1
                for (int c0 = 0; c0 < 32; c0 += U) {
                    float xr[U];
#pragma unroll
                    for (int i = 0; i < U; i++) xr[i] = sx[c0 + i];
#pragma unroll
                    for (int i = 0; i < U; i++) acc += biquad_step_fast(xr[i], m0, m1);
                }
                sy[0] = acc;
            }
        } else if (V == 7) {                // chain + y stores + per-sample state stores
            if (lane == 0) {
                float xr[32];
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = sx[i];
#pragma unroll
                for (int i = 0; i < 32; i++) { sm0[i] = m0; sm1[i] = m1; sy[i] = biquad_step_fast(xr[i], m0, m1); }
                sm0[32] = m0; sm1[32] = m1;
            }
        } else if (V == 8 || V == 9) {      // V8: fast pass + trivial phases; V9: + verification but no tiny/flag logic
            float e0 = 0, e1 = 0;
            if (lane == 0) {
                float xr[32];
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = sx[i];
                e0 = m0; e1 = m1;
#pragma unroll
                for (int i = 0; i < 32; i++) { sm0[i] = e0; sm1[i] = e1; sy[i] = biquad_step_fast(xr[i], e0, e1); }
                sm0[32] = e0; sm1[32] = e1;
            }
            __syncwarp();
            bool hz = false;
            if (V == 9) {
                float a = sm0[lane], b2 = sm1[lane];
                const float yv = biquad_step_eft(sx[lane], a, b2, hz);
                hz |= (f2u(a) != f2u(sm0[lane + 1])) | (f2u(b2) != f2u(sm1[lane + 1])) | (f2u(yv) != f2u(sy[lane]));
            }
            const unsigned any = __ballot_sync(0xffffffffu, hz);
            if (lane == 0) { if (!any) { m0 = e0; m1 = e1; } else { for (int i = 0; i < 32; i++) sy[i] = biquad_step(sx[i], m0, m1); } }
        } else if (V == 3) {                // double path
            if (lane == 0) {
#pragma unroll 8
                for (int i = 0; i < 32; i++) sy[i] = biquad_step(sx[i], m0, m1);
            }
        }
        __syncwarp();
        y[b * 32 + lane] = sy[lane];
    }
    if (lane == 0) *cyc = clock64() - t0;
}
// two-warp structure of k_rn_biquad: warp 0 lane 0 chain on block g, warp 1 verifies block g-1; W = what warp 1 does
template <int W> __global__ void __launch_bounds__(64) run2(const float *__restrict__ x, float *__restrict__ y, long long *cyc) {
    __shared__ float sx[2][32], sy[2][32], P0[2][33], P1[2][33];
    __shared__ int F[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float m0 = 0, m1 = 0;
    if (threadIdx.x < 2) F[threadIdx.x] = 0;
    if (warp == 1) sx[0][lane] = x[lane];
    __syncthreads();
    long long t0 = clock64();
    for (int g = 0; g <= NB; g++) {
        if (warp == 0) {
            if (lane == 0 && g < NB) {
                const int par = g & 1;
                float xr[32];
#pragma unroll
                for (int i = 0; i < 32; i++) xr[i] = sx[par][i];
#pragma unroll
                for (int i = 0; i < 32; i++) { P0[par][i] = m0; P1[par][i] = m1; sy[par][i] = biquad_step_fast(xr[i], m0, m1); }
                P0[par][32] = m0; P1[par][32] = m1;
            }
        } else {
            if (g >= 1) {
                const int par = (g - 1) & 1;
                bool hz = false;
                if (W >= 1) {
                    float a = P0[par][lane], b = P1[par][lane];
                    const float yv = biquad_step_eft(sx[par][lane], a, b, hz);
                    hz |= (f2u(a) != f2u(P0[par][lane + 1])) | (f2u(b) != f2u(P1[par][lane + 1])) | (f2u(yv) != f2u(sy[par][lane]));
                }
                const unsigned any = __ballot_sync(0xffffffffu, hz);
                if (lane == 0) F[par] = any != 0u;
                y[(g - 1) * 32 + lane] = sy[par][lane];
            }
            if (W >= 2 && g + 1 < NB) sx[(g + 1) & 1][lane] = x[(g + 1) * 32 + lane];
        }
        __syncthreads();
        if (threadIdx.x == 0 && g >= 1 && F[(g - 1) & 1]) {
            const int par = (g - 1) & 1;
            float r0 = P0[par][0], r1 = P1[par][0];
            for (int i = 0; i < 32; i++) sy[par][i] = biquad_step(sx[par][i], r0, r1);
        }
        if (W != 3) __syncthreads();
    }
    if (threadIdx.x == 0) *cyc = clock64() - t0;
}
int main() {
    const int N = NB * 32;
    float *hx = (float *)malloc(N * 4), *x, *y; long long *cyc, c;
    unsigned s = 12345; for (int i = 0; i < N; i++) { s = s * 1664525u + 1013904223u; hx[i] = ((int)(s >> 8) - (1 << 23)) / (float)(1 << 24) * 0.06f; }
    cudaMalloc(&x, N * 4); cudaMalloc(&y, N * 4); cudaMalloc(&cyc, 8);
    cudaMemcpy(x, hx, N * 4, cudaMemcpyHostToDevice);
    const char *nm[] = {"fast chain, regs only", "fast chain + y stores", "verified block (kernel path)", "double path", "fast chain unroll 4", "fast chain unroll 8", "fast chain unroll 16", "chain + y + state stores", "fast + ballot, no verify", "fast + verify + ballot"};
    for (int rep = 0; rep < 2; rep++) {
#define R(V) run<V><<<1, 32>>>(x, y, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("%-30s: %.1f cyc/sample\n", nm[V], (double)c / N);
        R(0) R(1) R(2) R(3) R(4) R(5) R(6) R(7) R(8) R(9)
    }
    for (int rep = 0; rep < 2; rep++) {
#define R2(W, name) run2<W><<<1, 64>>>(x, y, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("%-30s: %.1f cyc/sample\n", name, (double)c / N);
        R2(0, "2 warps: barriers only") R2(1, "2 warps: + verify") R2(2, "2 warps: + verify + loads") R2(3, "2 warps: one barrier")
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
