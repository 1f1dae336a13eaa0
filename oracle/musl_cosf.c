/*
 * musl_cosf.c -- restatement of musl libc's cosf (src/math/cosf.c, __cosdf.c, __sindf.c; FreeBSD msun s_cosf.c, k_cosf.c,
 * k_sinf.c), the algorithm Zig's compiler-rt ports for `@cos` on f32 (lib/compiler_rt/cos.zig: "Ported from musl").
 * TEST INFRASTRUCTURE ONLY.  Third-party algorithm absent from /root/reference: the reference's Hann window
 * (src/audio_utils/window_fn.zig:51-68) calls `@cos` on f32 arguments in [0, 2 pi); this file pins what that call
 * returns so that the window table the oracle and the CUDA path use ((float)cos((double)x), glibc) can be compared with it
 * entry by entry (tests/test_oracle.py::test_hann_window_cosf_pin).  Only the branches |x| <= 9 pi / 4 are needed for that;
 * larger arguments would need __rem_pio2f and return NaN here.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* |cos(x) - c(x)| < 2**-34.1 (~[-5.37e-11, 5.295e-11]). */
static const double C0 = -0x1ffffffd0c5e81.0p-54, /* -0.499999997251031003120 */
                    C1 = 0x155553e1053a42.0p-57,  /*  0.0416666233237390631894 */
                    C2 = -0x16c087e80f1e27.0p-62, /* -0.00138867637746099294692 */
                    C3 = 0x199342e0ee5069.0p-68;  /*  0.0000243904487962774090654 */
static float cosdf(double x) {
    double r, w, z;
    z = x * x;
    w = z * z;
    r = C2 + z * C3;
    return (float)(((1.0 + z * C0) + w * C1) + (w * z) * r);
}
/* |sin(x)/x - s(x)| < 2**-37.5 (~[-4.89e-12, 4.824e-12]). */
static const double S1 = -0x15555554cbac77.0p-55, /* -0.166666666416265235595 */
                    S2 = 0x111110896efbb2.0p-59,  /*  0.0083333293858894631756 */
                    S3 = -0x1a00f9e2cae774.0p-65, /* -0.000198393348360966317347 */
                    S4 = 0x16cd878c3b46a7.0p-71;  /*  0.0000027183114939898219064 */
static float sindf(double x) {
    double r, s, w, z;
    z = x * x;
    w = z * z;
    r = S3 + z * S4;
    s = z * x;
    return (float)((x + s * (S1 + z * S2)) + s * w * r);
}

static const double c1pio2 = 1 * 1.57079632679489661923, c2pio2 = 2 * 1.57079632679489661923,
                    c3pio2 = 3 * 1.57079632679489661923, c4pio2 = 4 * 1.57079632679489661923;

float orc_musl_cosf(float x) {
    uint32_t ix;
    memcpy(&ix, &x, 4);
    const unsigned sign = ix >> 31;
    ix &= 0x7fffffff;
    if (ix <= 0x3f490fda) {               /* |x| ~<= pi/4 */
        if (ix < 0x39800000) return 1.0f; /* |x| < 2**-12 */
        return cosdf(x);
    }
    if (ix <= 0x407b53d1) {               /* |x| ~<= 5*pi/4 */
        if (ix > 0x4016cbe3)              /* |x|  ~> 3*pi/4 */
            return -cosdf(sign ? x + c2pio2 : x - c2pio2);
        return sign ? sindf(x + c1pio2) : sindf(c1pio2 - x);
    }
    if (ix <= 0x40e231d5) {               /* |x| ~<= 9*pi/4 */
        if (ix > 0x40afeddf)              /* |x| ~> 7*pi/4 */
            return cosdf(sign ? x + c4pio2 : x - c4pio2);
        return sign ? sindf(-x - c3pio2) : sindf(x - c3pio2);
    }
    return NAN;                           /* needs __rem_pio2f: outside what the Hann window asks for */
}

/* window_fn.hannWindowPeriodic (window_fn.zig:22-28,51-68) with musl's cosf in place of (float)cos((double)x). */
void orc_hann_window_periodic_musl(float *w, size_t n) {
    const float N = (float)n;
    const float two_pi = (float)(2 * 3.14159265358979323846264338327950288419716939937510);
    for (size_t i = 0; i < n; i++) {
        const float nf = (float)i;
        float r = 0;
        r += (1.0f * 0.5f) * orc_musl_cosf(((two_pi * 0.0f) * nf) / N);
        r += (-1.0f * 0.5f) * orc_musl_cosf(((two_pi * 1.0f) * nf) / N);
        w[i] = r;
    }
}
