// What is the SM clock while a few SMs run a dependent chain?  clock64() against %globaltimer.  (round 2)
#include <cstdio>
__global__ void k(float *out, long long *res, int iters) {
    float a = out[threadIdx.x & 31], b = 1.0001f;
    unsigned long long g0, g1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    const long long c0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 64; k++) a = a * b + 0.5f;
    }
    const long long c1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0) { res[2 * blockIdx.x] = c1 - c0; res[2 * blockIdx.x + 1] = (long long)(g1 - g0); }
}
int main() {
    float *out; long long *res, h[2 * 148];
    cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&res, sizeof h);
    cudaMemset(out, 0, 148 * 128 * 4);
    for (int grid : {1, 20, 148}) for (int threads : {32, 128}) for (int rep = 0; rep < 2; rep++) {
        k<<<grid, threads>>>(out, res, 400000);
        cudaMemcpy(h, res, sizeof h, cudaMemcpyDeviceToHost);
        if (rep) printf("grid %3d x %3d threads: %.1f ms, SM clock %.1f MHz (block 0), %.1f MHz (last block)\n", grid, threads, h[1] / 1e6,
                        1e3 * h[0] / h[1], 1e3 * h[2 * (grid - 1)] / h[2 * (grid - 1) + 1]);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
