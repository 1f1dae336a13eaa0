/*
 * synth_host.c -- host-only build of the seeded input generator (include/fvad_synth_core.h) for the CPU arms of
 * bench.py (cpu_baseline, --impl reference) and for tests: TEST INFRASTRUCTURE ONLY.  It lets those arms generate
 * the very streams the GPU arm processes without mapping libfvad_cuda.so.  tests/test_oracle.py asserts it is
 * bit-equal to fvad_synth_fill_host.
 */
#include <math.h>
#include <stddef.h>
#include "../include/fvad_synth_core.h"

static float g_tab[FVAD_SYNTH_SIN_N];
static int g_ready;

static const float *tab(void) {
    if (!g_ready) {        /* idempotent: concurrent first calls write the same values */
        for (int i = 0; i < FVAD_SYNTH_SIN_N; i++) g_tab[i] = (float)sin(FVAD_SYNTH_TWO_PI * i / FVAD_SYNTH_SIN_N);
        __atomic_store_n(&g_ready, 1, __ATOMIC_RELEASE);
    }
    return g_tab;
}

int orc_synth_fill(float *dst, uint64_t seed, uint64_t stream_id, uint64_t first_sample, size_t n) {
    const float *t = tab();
    for (size_t i = 0; i < n; i++) dst[i] = fvad_synth_sample(t, seed, stream_id, first_sample + i);
    return 0;
}

/* Ly = log10(1e-2 + Ex) as the reference's C evaluates it (lib/rnnoise/src/denoise.c:349): the HOST libm's double log10.
 * Checker for tests/test_gpu_denoiser.py::test_device_log10_equals_host_log10_on_a_sweep. */
void orc_log10_ex(const float *ex, float *out, size_t n) {
    for (size_t i = 0; i < n; i++) out[i] = (float)log10(1e-2 + ex[i]);
}
