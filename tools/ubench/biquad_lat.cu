// Microbenchmark: latency of the exact biquad recurrence (denoise.c:410-420) in several equivalent
// formulations, and of the primitive dependent chains involved.  One thread, clock64() timing.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#define N 48000
__device__ __forceinline__ float step0(float xi, float &m0, float &m1) {
    const float a0 = -1.99599f, a1 = 0.99600f, b0 = -2.f, b1 = 1.f;
    const float yi = xi + m0;
    m0 = (float)(m1 + (b0 * (double)xi - a0 * (double)yi));
    m1 = (float)(b1 * (double)xi - a1 * (double)yi);
    return yi;
}
// DFMA formulation: products of two floats are exact in double, so DMUL+DADD == DFMA bit for bit.
__device__ __forceinline__ float step1(float xi, float &m0, float &m1) {
    const double A0 = (double)1.99599f, A1 = (double)0.99600f;
    const float yi = xi + m0;
    const double yd = (double)yi, xd = (double)xi;
    const double u = fma(A0, yd, -2.0 * xd);
    m0 = (float)((double)m1 + u);
    m1 = (float)fma(-A1, yd, xd);
    return yi;
}
__device__ __forceinline__ double f2d_bits(float f) {     // exact f32->f64 for normal numbers and zero via integer ops
    const unsigned u = __float_as_uint(f);
    const unsigned a = u & 0x7fffffffu;
    unsigned hi = (u & 0x80000000u) | ((a >> 3) + 0x38000000u);
    unsigned lo = u << 29;
    if (a < 0x00800000u) return (double)f;                   // zero / denormal: slow path
    return __hiloint2double((int)hi, (int)lo);
}
__device__ __forceinline__ float step2(float xi, float &m0, float &m1) {
    const double A0 = (double)1.99599f, A1 = (double)0.99600f;
    const float yi = xi + m0;
    const double yd = f2d_bits(yi), xd = (double)xi;
    const double u = fma(A0, yd, -2.0 * xd);
    m0 = (float)(f2d_bits(m1) + u);
    m1 = (float)fma(-A1, yd, xd);
    return yi;
}
__device__ __forceinline__ float step3(float xi, float &m0, float &m1) {   // NOT exact: float FMA floor
    const float yi = xi + m0;
    const float c = m1 - 2.f * xi;
    m0 = __fmaf_rn(1.99599f, yi, c);
    m1 = __fmaf_rn(-0.99600f, yi, xi);
    return yi;
}
template <int V> __global__ void run(const float *x, float *y, long long *cyc) {
    float m0 = 0, m1 = 0;
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N; i++) {
        float xi = x[i] * 32767.0f;
        y[i] = V == 0 ? step0(xi, m0, m1) : V == 1 ? step1(xi, m0, m1) : V == 2 ? step2(xi, m0, m1) : step3(xi, m0, m1);
    }
    *cyc = clock64() - t0;
}
template <int K> __global__ void prim(double seed, float fseed, long long *cyc, double *sink) {
    double d = seed; float f = fseed;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < 4096; i++) {
        if (K == 0) d = d + 1.0000001;                         // DADD chain
        if (K == 1) d = fma(d, 0.9999999, 1e-9);               // DFMA chain
        if (K == 2) { f = (float)((double)f + 0.5); }          // F2D + DADD + D2F
        if (K == 3) { d = (double)(float)d; d = d * 1.0000001; } // D2F + F2D + DMUL
        if (K == 4) f = f + 1.0000001f;                        // FADD chain
        if (K == 5) { d = (double)(float)d; }                  // D2F + F2D only (compiler may keep)
        if (K == 6) d = d * 1.0000001;                         // DMUL chain
    }
    *cyc = clock64() - t0;
    *sink = d + f;
}
int main() {
    float *hx = (float *)malloc(N * 4), *x, *y; long long *cyc; double *sink;
    unsigned s = 12345; for (int i = 0; i < N; i++) { s = s * 1664525u + 1013904223u; hx[i] = ((int)(s >> 8) - (1 << 23)) / (float)(1 << 24) * 0.3f; }
    cudaMalloc(&x, N * 4); cudaMalloc(&y, N * 4 * 4); cudaMalloc(&cyc, 8); cudaMalloc(&sink, 8);
    cudaMemcpy(x, hx, N * 4, cudaMemcpyHostToDevice);
    float *hy = (float *)malloc(N * 4 * 4); long long c;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    for (int rep = 0; rep < 2; rep++) {
        run<0><<<1, 1>>>(x, y, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("V0 exact ref : %.1f cyc/sample\n", (double)c / N);
        run<1><<<1, 1>>>(x, y + N, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("V1 dfma      : %.1f cyc/sample\n", (double)c / N);
        run<2><<<1, 1>>>(x, y + 2 * N, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("V2 dfma+bits : %.1f cyc/sample\n", (double)c / N);
        run<3><<<1, 1>>>(x, y + 3 * N, cyc); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); if (rep) printf("V3 float fma : %.1f cyc/sample (not exact)\n", (double)c / N);
    }
    cudaMemcpy(hy, y, N * 16, cudaMemcpyDeviceToHost);
    int d1 = 0, d2 = 0, d3 = 0;
    for (int i = 0; i < N; i++) { d1 += memcmp(&hy[i], &hy[N + i], 4) != 0; d2 += memcmp(&hy[i], &hy[2 * N + i], 4) != 0; d3 += memcmp(&hy[i], &hy[3 * N + i], 4) != 0; }
    printf("mismatches vs V0: V1 %d  V2 %d  V3 %d of %d\n", d1, d2, d3, N);
    const char *names[] = {"DADD", "DFMA", "F2D+DADD+D2F", "D2F+F2D+DMUL", "FADD", "D2F+F2D", "DMUL"};
#define P(K) prim<K><<<1, 1>>>(1.0, 1.0f, cyc, sink); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); printf("%-14s: %.1f cyc/iter\n", names[K], (double)c / 4096);
    P(0) P(1) P(2) P(3) P(4) P(5) P(6)
    printf("clock %d kHz\n", clk);
    return 0;
}
