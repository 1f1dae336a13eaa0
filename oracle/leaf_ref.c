/*
 * leaf_ref.c -- binds the oracle's leaf seam to the REFERENCE's own C libraries,
 * compiled unmodified from /root/reference by oracle/Makefile (target `ref`):
 *   lib/kissfft/{kiss_fft.c,kiss_fftr.c}  (-Dkiss_fft_scalar=float, build.zig:145-147)
 *   lib/rnnoise/src/{denoise,rnn,rnn_data,pitch,celt_lpc,kiss_fft}.c (build.zig:183-198)
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  No reference source is copied here.
 */
#include "leaf.h"
#include <stdlib.h>
#include <string.h>
#include "kiss_fftr.h"     /* -I/root/reference/lib/kissfft */
#include "rnnoise.h"       /* -I/root/reference/lib/rnnoise/include */

struct leaf_rfft { int nfft; void *mem; size_t len; kiss_fftr_cfg cfg; };

leaf_rfft *leaf_rfft_create(int nfft) {
    if (nfft <= 0 || (nfft & 1)) return NULL;
    leaf_rfft *f = (leaf_rfft *)calloc(1, sizeof(*f));
    f->nfft = nfft;
    size_t len = 1;
    kiss_fftr_alloc(nfft, 0, NULL, &len);          /* FFT.zig:193-208 kissfftSizeNeeded */
    f->len = len;
    f->mem = aligned_alloc(16, (len + 15) & ~(size_t)15);
    leaf_rfft_reset(f);
    return f;
}
void leaf_rfft_reset(leaf_rfft *f) {               /* FFT.zig:179-191 resetKissFFT */
    size_t len = f->len;
    f->cfg = kiss_fftr_alloc(f->nfft, 0, f->mem, &len);
}
void leaf_rfft_run(leaf_rfft *f, const float *in, float *out_cpx) {
    kiss_fftr(f->cfg, in, (kiss_fft_cpx *)out_cpx); /* FFT.zig:90 */
}
void leaf_rfft_destroy(leaf_rfft *f) { if (f) { free(f->mem); free(f); } }

struct leaf_rnn { DenoiseState *st; };
leaf_rnn *leaf_rnn_create(void) {
    leaf_rnn *r = (leaf_rnn *)calloc(1, sizeof(*r));
    r->st = rnnoise_create(NULL);                  /* Denoiser.zig:23 */
    return r;
}
void leaf_rnn_destroy(leaf_rnn *r) { if (r) { rnnoise_destroy(r->st); free(r); } }
float leaf_rnn_process(leaf_rnn *r, float *out, const float *in, orc_rnnoise_trace *tr) {
    float vad = rnnoise_process_frame(r->st, out, in);   /* Denoiser.zig:60 */
    if (tr) {                                      /* internals are not reachable through the library API */
        memset(tr, 0, sizeof(*tr));
        tr->pitch_index = -1;
        tr->vad_prob = vad;
        memcpy(tr->out, out, sizeof(float) * 480);
    }
    return vad;
}
const char *leaf_kind(void) { return "reference"; }
