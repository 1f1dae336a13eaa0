/* leaf.h -- internal seam between the Zig-glue restatement (glue.c) and the
 * leaf arithmetic (leaf_port.c = restated C, leaf_ref.c = the reference's own
 * C libraries).  TEST INFRASTRUCTURE ONLY (see oracle.h). */
#ifndef FVAD_ORACLE_LEAF_H
#define FVAD_ORACLE_LEAF_H
#include "oracle.h"

typedef struct leaf_rfft leaf_rfft;
leaf_rfft *leaf_rfft_create(int nfft);           /* kiss_fftr_alloc equivalent */
void       leaf_rfft_reset(leaf_rfft *f);        /* regenerate state (FFT.zig:84 resetKissFFT) */
void       leaf_rfft_run(leaf_rfft *f, const float *in, float *out_cpx); /* nfft/2+1 (re,im) */
void       leaf_rfft_destroy(leaf_rfft *f);

typedef struct leaf_rnn leaf_rnn;
leaf_rnn *leaf_rnn_create(void);                 /* rnnoise_create(NULL) */
void      leaf_rnn_destroy(leaf_rnn *r);
float     leaf_rnn_process(leaf_rnn *r, float *out, const float *in, orc_rnnoise_trace *trace);

const char *leaf_kind(void);
#endif
