/*
 * leaf_port.c -- plain-C restatement of the reference's leaf arithmetic.
 * TEST INFRASTRUCTURE ONLY (see oracle.h): the parity checker and the "port"
 * CPU baseline.  Must be compiled with -ffp-contract=off (no FMA contraction):
 * every product and sum below is a separately rounded IEEE operation, exactly
 * as the reference's C is under `gcc -O2 -ffp-contract=off`.
 *
 * Part 1: KissFFT real FFT, float build (lib/kissfft/kiss_fft.c, kiss_fftr.c).
 * Part 2: RNNoise (lib/rnnoise/src/{denoise,rnn,pitch,celt_lpc,kiss_fft}.c).
 * The structure is ours (iterative stages, explicit tables); each function
 * cites the reference lines whose arithmetic and operation ORDER it follows.
 */
#include "leaf.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float r, i; } cpx;

static inline cpx c_mul(cpx a, cpx b) {          /* C_MUL, _kiss_fft_guts.h:87-89 */
    cpx m;
    m.r = a.r * b.r - a.i * b.i;
    m.i = a.r * b.i + a.i * b.r;
    return m;
}
static inline cpx c_add(cpx a, cpx b) { cpx m = { a.r + b.r, a.i + b.i }; return m; }
static inline cpx c_sub(cpx a, cpx b) { cpx m = { a.r - b.r, a.i - b.i }; return m; }

/* ======================================================================= */
/* Part 1: kiss_fftr for power-of-two nfft                                  */
/* ======================================================================= */
#define RF_MAXFAC 32
struct leaf_rfft {
    int nfft;            /* real length */
    int nc;              /* complex length nfft/2 */
    int nfac;
    int fac_p[RF_MAXFAC], fac_m[RF_MAXFAC];
    int *perm;           /* output position -> input index (the m==1 leaf copy, kiss_fft.c:274-278) */
    cpx *tw;             /* nc twiddles, kiss_fft.c:357-363 */
    cpx *stw;            /* nc/2 super twiddles, kiss_fftr.c:53-59 */
    cpx *tmp;
};

/* kf_factor (kiss_fft.c:307-328): 4s first, then 2s, then odd primes. */
static int rf_factor(int n, int *fp, int *fm) {
    int p = 4, k = 0;
    double floor_sqrt = floor(sqrt((double)n));
    do {
        while (n % p) {
            switch (p) { case 4: p = 2; break; case 2: p = 3; break; default: p += 2; break; }
            if (p > floor_sqrt) p = n;
        }
        n /= p;
        fp[k] = p; fm[k] = n; k++;
    } while (n > 1);
    return k;
}

/* The recursion of kf_work (kiss_fft.c:235-300) only decides WHICH input lands at
 * which output slot before the butterflies run; record that as a table. */
static void rf_perm(int *perm, int out_pos, int in_idx, int fstride, const int *fp, const int *fm, int lvl) {
    int p = fp[lvl], m = fm[lvl];
    for (int k = 0; k < p; k++) {
        if (m == 1) perm[out_pos + k] = in_idx + k * fstride;
        else rf_perm(perm, out_pos + k * m, in_idx + k * fstride, fstride * p, fp, fm, lvl + 1);
    }
}

void leaf_rfft_reset(leaf_rfft *f) {
    int nc = f->nc;
    for (int i = 0; i < nc; i++) {               /* kiss_fft.c:357-363 */
        const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
        double phase = -2 * pi * i / nc;
        f->tw[i].r = (float)cos(phase);
        f->tw[i].i = (float)sin(phase);
    }
    for (int i = 0; i < nc / 2; i++) {           /* kiss_fftr.c:53-59 */
        double phase = -3.14159265358979323846264338327 * ((double)(i + 1) / nc + .5);
        f->stw[i].r = (float)cos(phase);
        f->stw[i].i = (float)sin(phase);
    }
    f->nfac = rf_factor(nc, f->fac_p, f->fac_m);
}

leaf_rfft *leaf_rfft_create(int nfft) {
    if (nfft < 4 || (nfft & (nfft - 1))) return NULL;   /* the port covers powers of two */
    leaf_rfft *f = (leaf_rfft *)calloc(1, sizeof(*f));
    f->nfft = nfft; f->nc = nfft / 2;
    f->perm = (int *)malloc(sizeof(int) * f->nc);
    f->tw = (cpx *)malloc(sizeof(cpx) * f->nc);
    f->stw = (cpx *)malloc(sizeof(cpx) * (f->nc / 2 + 1));
    f->tmp = (cpx *)malloc(sizeof(cpx) * f->nc);
    leaf_rfft_reset(f);
    rf_perm(f->perm, 0, 0, 1, f->fac_p, f->fac_m, 0);
    return f;
}

void leaf_rfft_destroy(leaf_rfft *f) {
    if (!f) return;
    free(f->perm); free(f->tw); free(f->stw); free(f->tmp); free(f);
}

/* kf_bfly2 (kiss_fft.c:15-36) */
static void rf_bfly2(cpx *F, int fstride, const cpx *tw, int m) {
    for (int k = 0; k < m; k++) {
        cpx t = c_mul(F[m + k], tw[k * fstride]);
        F[m + k] = c_sub(F[k], t);
        F[k] = c_add(F[k], t);
    }
}

/* kf_bfly4, forward branch (kiss_fft.c:38-84) */
static void rf_bfly4(cpx *F, int fstride, const cpx *tw, int m) {
    for (int k = 0; k < m; k++) {
        cpx s0 = c_mul(F[k + m], tw[k * fstride]);
        cpx s1 = c_mul(F[k + 2 * m], tw[k * fstride * 2]);
        cpx s2 = c_mul(F[k + 3 * m], tw[k * fstride * 3]);
        cpx s5 = c_sub(F[k], s1);
        F[k] = c_add(F[k], s1);
        cpx s3 = c_add(s0, s2);
        cpx s4 = c_sub(s0, s2);
        F[k + 2 * m] = c_sub(F[k], s3);
        F[k] = c_add(F[k], s3);
        F[k + m].r = s5.r + s4.i;
        F[k + m].i = s5.i - s4.r;
        F[k + 3 * m].r = s5.r - s4.i;
        F[k + 3 * m].i = s5.i + s4.r;
    }
}

void leaf_rfft_run(leaf_rfft *f, const float *in, float *out_cpx) {
    const int nc = f->nc;
    const cpx *fin = (const cpx *)in;            /* kiss_fftr.c:77: reals packed as complex */
    cpx *t = f->tmp;
    cpx *out = (cpx *)out_cpx;
    for (int i = 0; i < nc; i++) t[i] = fin[f->perm[i]];
    /* deepest factor first; fstride at level L = product of the radices above it */
    int fstr[RF_MAXFAC + 1];
    fstr[0] = 1;
    for (int L = 0; L < f->nfac; L++) fstr[L + 1] = fstr[L] * f->fac_p[L];
    for (int L = f->nfac - 1; L >= 0; L--) {
        int p = f->fac_p[L], m = f->fac_m[L];
        for (int base = 0; base < nc; base += p * m) {
            if (p == 4) rf_bfly4(t + base, fstr[L], f->tw, m);
            else rf_bfly2(t + base, fstr[L], f->tw, m);
        }
    }
    /* kiss_fftr.c:90-116 */
    float tr = t[0].r, ti = t[0].i;
    out[0].r = tr + ti;
    out[nc].r = tr - ti;
    out[0].i = 0; out[nc].i = 0;
    for (int k = 1; k <= nc / 2; k++) {
        cpx fpk = t[k];
        cpx fpnk = { t[nc - k].r, -t[nc - k].i };
        cpx f1k = c_add(fpk, fpnk);
        cpx f2k = c_sub(fpk, fpnk);
        cpx w = c_mul(f2k, f->stw[k - 1]);
        out[k].r = (f1k.r + w.r) * 0.5f;
        out[k].i = (f1k.i + w.i) * 0.5f;
        out[nc - k].r = (f1k.r - w.r) * 0.5f;
        out[nc - k].i = (w.i - f1k.i) * 0.5f;
    }
}

/* ======================================================================= */
/* Part 2: RNNoise                                                          */
/* ======================================================================= */
#define FRAME 480
#define WIN 960
#define NFREQ 481
#define NB 22
#define PBUF 1728            /* PITCH_BUF_SIZE, denoise.c:52-55 */
#define PMAX 768
#define PMIN 60
#define CEPS 8
#define NDELTA 6
#define NFEAT 42

static const int eband[NB] = { 0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 34, 40, 48, 60, 78, 100 }; /* denoise.c:73-76 */

/* ---- model blob (formula-vad_b200/data/rnnoise_model.bin; tools/extract_rnnoise_model.py) */
#define _GNU_SOURCE_DL 1
#include <dlfcn.h>
typedef struct {
    const signed char *in_w, *in_b;                         /* dense 42->24 tanh      rnn_data.c:145-149 */
    const signed char *vg_w, *vg_r, *vg_b;                  /* GRU 24->24 relu        rnn_data.c:601-606 */
    const signed char *ng_w, *ng_r, *ng_b;                  /* GRU 90->48 relu        rnn_data.c:3119-3124 */
    const signed char *dg_w, *dg_r, *dg_b;                  /* GRU 114->96 relu       rnn_data.c:10731-10736 */
    const signed char *do_w, *do_b;                         /* dense 96->22 sigmoid   rnn_data.c:11011-11015 */
    const signed char *vo_w, *vo_b;                         /* dense 24->1 sigmoid    rnn_data.c:11027-11031 */
    const float *tansig;                                    /* tansig_table.h:3 */
} rn_model;
static rn_model g_model;
static unsigned char *g_blob;

static const void *blob_find(const char *name, unsigned want_count) {
    unsigned n = *(unsigned *)(g_blob + 8);
    size_t off = 12 + (size_t)n * 40;
    for (unsigned a = 0; a < n; a++) {
        const unsigned char *e = g_blob + 12 + (size_t)a * 40;
        unsigned dt = *(unsigned *)(e + 32), cnt = *(unsigned *)(e + 36);
        size_t bytes = (size_t)cnt * (dt ? 4 : 1);
        if (strncmp((const char *)e, name, 32) == 0) {
            if (cnt != want_count) { fprintf(stderr, "oracle: blob array %s has %u, want %u\n", name, cnt, want_count); abort(); }
            return g_blob + off;
        }
        off += (bytes + 15) & ~(size_t)15;
    }
    fprintf(stderr, "oracle: blob array %s missing\n", name); abort();
}

static void model_load(void) {
    if (g_blob) return;
    char path[4096];
    const char *env = getenv("FVAD_MODEL_BLOB");
    if (env) snprintf(path, sizeof path, "%s", env);
    else {
        Dl_info info;
        if (!dladdr((void *)&model_load, &info) || !info.dli_fname) { fprintf(stderr, "oracle: dladdr failed\n"); abort(); }
        snprintf(path, sizeof path, "%s", info.dli_fname);
        /* <repo>/oracle/liboracle_port.so or <repo>/oracle/_ref/liboracle_ref.so -> <repo> */
        char *s = strstr(path, "/oracle/");
        char *last = NULL;
        while (s) { last = s; s = strstr(s + 1, "/oracle/"); }
        if (!last) { fprintf(stderr, "oracle: cannot locate repo root from %s\n", path); abort(); }
        snprintf(last, sizeof path - (size_t)(last - path), "/formula-vad_b200/data/rnnoise_model.bin");
    }
    FILE *fp = fopen(path, "rb");
    if (!fp) { fprintf(stderr, "oracle: cannot open model blob %s\n", path); abort(); }
    fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    g_blob = (unsigned char *)malloc((size_t)sz);
    if (fread(g_blob, 1, (size_t)sz, fp) != (size_t)sz || memcmp(g_blob, "FVADRNN1", 8)) { fprintf(stderr, "oracle: bad blob\n"); abort(); }
    fclose(fp);
    g_model.in_w = blob_find("input_dense_weights", 1008);  g_model.in_b = blob_find("input_dense_bias", 24);
    g_model.vg_w = blob_find("vad_gru_weights", 1728);      g_model.vg_r = blob_find("vad_gru_recurrent_weights", 1728);
    g_model.vg_b = blob_find("vad_gru_bias", 72);
    g_model.ng_w = blob_find("noise_gru_weights", 12960);   g_model.ng_r = blob_find("noise_gru_recurrent_weights", 6912);
    g_model.ng_b = blob_find("noise_gru_bias", 144);
    g_model.dg_w = blob_find("denoise_gru_weights", 32832); g_model.dg_r = blob_find("denoise_gru_recurrent_weights", 27648);
    g_model.dg_b = blob_find("denoise_gru_bias", 288);
    g_model.do_w = blob_find("denoise_output_weights", 2112); g_model.do_b = blob_find("denoise_output_bias", 22);
    g_model.vo_w = blob_find("vad_output_weights", 24);     g_model.vo_b = blob_find("vad_output_bias", 1);
    g_model.tansig = blob_find("tansig_table", 201);
}

/* ---- tables: check_init (denoise.c:168-182), opus_fft_alloc_twiddles (kiss_fft.c:434-493) */
static int   g_tab_init;
static float g_half_window[FRAME];
static float g_dct[NB * NB];
static cpx   g_tw960[WIN];
static short g_rev960[WIN];

static void tables_init(void) {
    if (g_tab_init) return;
    for (int i = 0; i < FRAME; i++)
        g_half_window[i] = (float)sin(.5 * M_PI * sin(.5 * M_PI * (i + .5) / FRAME) * sin(.5 * M_PI * (i + .5) / FRAME));
    for (int i = 0; i < NB; i++)
        for (int j = 0; j < NB; j++) {
            g_dct[i * NB + j] = (float)cos((i + .5) * j * M_PI / NB);
            if (j == 0) g_dct[i * NB + j] = (float)(g_dct[i * NB + j] * sqrt(.5));
        }
    for (int i = 0; i < WIN; i++) {              /* compute_twiddles, kiss_fft.c:404-419 */
        const double pi = 3.14159265358979323846264338327;
        double phase = (-2 * pi / WIN) * i;
        g_tw960[i].r = (float)cos(phase);
        g_tw960[i].i = (float)sin(phase);
    }
    /* factors after kf_factor's reversal (kiss_fft.c:351-401): radix 5,3,4,4,4 with
     * m = 192,64,16,4,1.  compute_bitrev_table (kiss_fft.c:314-343) then sends input
     * n = n0 + 5 n1 + 15 n2 + 60 n3 + 240 n4 to slot 192 n0 + 64 n1 + 16 n2 + 4 n3 + n4. */
    for (int n = 0; n < WIN; n++) {
        int n0 = n % 5, n1 = (n / 5) % 3, n2 = (n / 15) % 4, n3 = (n / 60) % 4, n4 = n / 240;
        g_rev960[n] = (short)(192 * n0 + 64 * n1 + 16 * n2 + 4 * n3 + n4);
    }
    g_tab_init = 1;
}

/* ---- the 960-point Opus FFT: opus_fft_c + opus_fft_impl (kiss_fft.c:518-586) */
static void fft960_r4_m1(cpx *F) {               /* kf_bfly4 m==1 branch, kiss_fft.c:112-131 */
    for (int i = 0; i < 240; i++, F += 4) {
        cpx s0 = c_sub(F[0], F[2]);
        F[0] = c_add(F[0], F[2]);
        cpx s1 = c_add(F[1], F[3]);
        F[2] = c_sub(F[0], s1);
        F[0] = c_add(F[0], s1);
        s1 = c_sub(F[1], F[3]);
        F[1].r = s0.r + s1.i;
        F[1].i = s0.i - s1.r;
        F[3].r = s0.r - s1.i;
        F[3].i = s0.i + s1.r;
    }
}
static void fft960_r4(cpx *Fb, int fstride, int m, int N, int mm) {   /* kiss_fft.c:132-168 */
    for (int i = 0; i < N; i++) {
        cpx *F = Fb + i * mm;
        for (int j = 0; j < m; j++, F++) {
            cpx s0 = c_mul(F[m], g_tw960[j * fstride]);
            cpx s1 = c_mul(F[2 * m], g_tw960[j * fstride * 2]);
            cpx s2 = c_mul(F[3 * m], g_tw960[j * fstride * 3]);
            cpx s5 = c_sub(F[0], s1);
            F[0] = c_add(F[0], s1);
            cpx s3 = c_add(s0, s2);
            cpx s4 = c_sub(s0, s2);
            F[2 * m] = c_sub(F[0], s3);
            F[0] = c_add(F[0], s3);
            F[m].r = s5.r + s4.i;
            F[m].i = s5.i - s4.r;
            F[3 * m].r = s5.r - s4.i;
            F[3 * m].i = s5.i + s4.r;
        }
    }
}
static void fft960_r3(cpx *Fb, int fstride, int m, int N, int mm) {   /* kf_bfly3, kiss_fft.c:174-226 */
    const float epi3_i = g_tw960[fstride * m].i;
    for (int i = 0; i < N; i++) {
        cpx *F = Fb + i * mm;
        for (int k = 0; k < m; k++, F++) {
            cpx s1 = c_mul(F[m], g_tw960[k * fstride]);
            cpx s2 = c_mul(F[2 * m], g_tw960[k * fstride * 2]);
            cpx s3 = c_add(s1, s2);
            cpx s0 = c_sub(s1, s2);
            F[m].r = F[0].r - s3.r * .5f;
            F[m].i = F[0].i - s3.i * .5f;
            s0.r *= epi3_i;
            s0.i *= epi3_i;
            F[0] = c_add(F[0], s3);
            F[2 * m].r = F[m].r + s0.i;
            F[2 * m].i = F[m].i - s0.r;
            F[m].r = F[m].r - s0.i;
            F[m].i = F[m].i + s0.r;
        }
    }
}
static void fft960_r5(cpx *Fb, int fstride, int m, int N, int mm) {   /* kf_bfly5, kiss_fft.c:230-305 */
    const cpx ya = g_tw960[fstride * m], yb = g_tw960[fstride * 2 * m];
    for (int i = 0; i < N; i++) {
        cpx *F0 = Fb + i * mm, *F1 = F0 + m, *F2 = F0 + 2 * m, *F3 = F0 + 3 * m, *F4 = F0 + 4 * m;
        for (int u = 0; u < m; u++, F0++, F1++, F2++, F3++, F4++) {
            cpx s0 = *F0;
            cpx s1 = c_mul(*F1, g_tw960[u * fstride]);
            cpx s2 = c_mul(*F2, g_tw960[2 * u * fstride]);
            cpx s3 = c_mul(*F3, g_tw960[3 * u * fstride]);
            cpx s4 = c_mul(*F4, g_tw960[4 * u * fstride]);
            cpx s7 = c_add(s1, s4), s10 = c_sub(s1, s4), s8 = c_add(s2, s3), s9 = c_sub(s2, s3);
            F0->r = F0->r + (s7.r + s8.r);
            F0->i = F0->i + (s7.i + s8.i);
            cpx s5, s6, s11, s12;
            s5.r = s0.r + (s7.r * ya.r + s8.r * yb.r);
            s5.i = s0.i + (s7.i * ya.r + s8.i * yb.r);
            s6.r = s10.i * ya.i + s9.i * yb.i;
            s6.i = -(s10.r * ya.i + s9.r * yb.i);
            *F1 = c_sub(s5, s6);
            *F4 = c_add(s5, s6);
            s11.r = s0.r + (s7.r * yb.r + s8.r * ya.r);
            s11.i = s0.i + (s7.i * yb.r + s8.i * ya.r);
            s12.r = s9.i * ya.i - s10.i * yb.i;
            s12.i = s10.r * yb.i - s9.r * ya.i;
            *F2 = c_add(s11, s12);
            *F3 = c_sub(s11, s12);
        }
    }
}
static void fft960(const cpx *in, cpx *out) {
    const float scale = 1.f / WIN;                /* kiss_fft.c:459 */
    for (int i = 0; i < WIN; i++) {               /* kiss_fft.c:579-584 */
        out[g_rev960[i]].r = scale * in[i].r;
        out[g_rev960[i]].i = scale * in[i].i;
    }
    fft960_r4_m1(out);                            /* stage order, kiss_fft.c:539-563 */
    fft960_r4(out, 60, 4, 60, 16);
    fft960_r4(out, 15, 16, 15, 64);
    fft960_r3(out, 5, 64, 5, 192);
    fft960_r5(out, 1, 192, 1, 1);
}

/* ---- band helpers -------------------------------------------------------- */
/* compute_band_energy / compute_band_corr (denoise.c:100-148): triangular bands,
 * sequential f32 accumulation in (band, j) order; corr==NULL means energy of X. */
static void band_accumulate(float *bandE, const cpx *X, const cpx *P) {
    float sum[NB] = { 0 };
    for (int i = 0; i < NB - 1; i++) {
        int band_size = (eband[i + 1] - eband[i]) << 2;
        for (int j = 0; j < band_size; j++) {
            int b = (eband[i] << 2) + j;
            float frac = (float)j / band_size;
            float tmp;
            if (P) { tmp = X[b].r * P[b].r; tmp += X[b].i * P[b].i; }
            else   { tmp = X[b].r * X[b].r; tmp += X[b].i * X[b].i; }
            sum[i] += (1 - frac) * tmp;
            sum[i + 1] += frac * tmp;
        }
    }
    sum[0] *= 2;
    sum[NB - 1] *= 2;
    for (int i = 0; i < NB; i++) bandE[i] = sum[i];
}

/* interp_band_gain (denoise.c:150-163).  The reference's memset clears 481 BYTES
 * (not floats); callers pre-zero the arrays, so the net effect is: bins 0..399
 * interpolated, bins 400..480 keep the caller's initial value (0). */
static void band_interp(float *g, const float *bandE) {
    for (int i = 0; i < NB - 1; i++) {
        int band_size = (eband[i + 1] - eband[i]) << 2;
        for (int j = 0; j < band_size; j++) {
            float frac = (float)j / band_size;
            g[(eband[i] << 2) + j] = (1 - frac) * bandE[i] + frac * bandE[i + 1];
        }
    }
}

static void dct22(float *out, const float *in) {  /* dct, denoise.c:184-195 */
    for (int i = 0; i < NB; i++) {
        float sum = 0;
        for (int j = 0; j < NB; j++) sum += in[j] * g_dct[j * NB + i];
        out[i] = (float)(sum * sqrt(2. / 22));
    }
}

static void window960(float *x) {                 /* apply_window, denoise.c:247-254 */
    for (int i = 0; i < FRAME; i++) {
        x[i] *= g_half_window[i];
        x[WIN - 1 - i] *= g_half_window[i];
    }
}

static void forward960(cpx *out481, const float *in) {   /* forward_transform, denoise.c:212-225 */
    cpx x[WIN], y[WIN];
    for (int i = 0; i < WIN; i++) { x[i].r = in[i]; x[i].i = 0; }
    fft960(x, y);
    for (int i = 0; i < NFREQ; i++) out481[i] = y[i];
}

/* ---- pitch analysis (pitch.c, celt_lpc.c) --------------------------------- */
static float inner(const float *x, const float *y, int n) {   /* celt_inner_prod, pitch.h:137-144 */
    float xy = 0;
    for (int i = 0; i < n; i++) xy = xy + x[i] * y[i];
    return xy;
}

/* pitch_downsample (pitch.c:148-216) with _celt_autocorr (celt_lpc.c:198-279),
 * _celt_lpc (celt_lpc.c:37-88) and celt_fir5 (pitch.c:106-145) for C==1. */
static void pitch_down(const float *x, float *x_lp, int len) {
    int n = len >> 1;
    for (int i = 1; i < n; i++)
        x_lp[i] = .5f * (.5f * (x[2 * i - 1] + x[2 * i + 1]) + x[2 * i]);
    x_lp[0] = .5f * (.5f * x[1] + x[0]);

    float ac[5];
    int lag = 4, fastN = n - lag;
    /* celt_pitch_xcorr(x,x,ac,fastN,5): each lag is its own in-order f32 chain
     * (xcorr_kernel, pitch.h:53-117, interleaves four lags but never reorders one). */
    for (int k = 0; k <= lag; k++) ac[k] = inner(x_lp, x_lp + k, fastN);
    for (int k = 0; k <= lag; k++) {
        float d = 0;
        for (int i = k + fastN; i < n; i++) d = d + x_lp[i] * x_lp[i - k];
        ac[k] += d;
    }
    ac[0] *= 1.0001f;
    for (int i = 1; i <= 4; i++) ac[i] -= ac[i] * (.008f * i) * (.008f * i);

    float lpc[4] = { 0, 0, 0, 0 };
    float error = ac[0];
    if (ac[0] != 0) {
        for (int i = 0; i < 4; i++) {
            float rr = 0;
            for (int j = 0; j < i; j++) rr += lpc[j] * ac[i - j];
            rr += ac[i + 1];
            float r = -rr / error;
            lpc[i] = r;
            for (int j = 0; j < (i + 1) >> 1; j++) {
                float t1 = lpc[j], t2 = lpc[i - 1 - j];
                lpc[j] = t1 + r * t2;
                lpc[i - 1 - j] = t2 + r * t1;
            }
            error = error - (r * r) * error;
            if (error < .001f * ac[0]) break;
        }
    }
    float tmp = 1.0f;
    for (int i = 0; i < 4; i++) { tmp = .9f * tmp; lpc[i] = lpc[i] * tmp; }
    const float c1 = .8f;
    float num[5];
    num[0] = lpc[0] + .8f;
    num[1] = lpc[1] + c1 * lpc[0];
    num[2] = lpc[2] + c1 * lpc[1];
    num[3] = lpc[3] + c1 * lpc[2];
    num[4] = c1 * lpc[3];
    float m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0;
    for (int i = 0; i < n; i++) {
        float xi = x_lp[i];
        float sum = xi;
        sum = sum + num[0] * m0;
        sum = sum + num[1] * m1;
        sum = sum + num[2] * m2;
        sum = sum + num[3] * m3;
        sum = sum + num[4] * m4;
        m4 = m3; m3 = m2; m2 = m1; m1 = m0; m0 = xi;
        x_lp[i] = sum;
    }
}

/* find_best_pitch, float build (pitch.c:46-104) */
static void best_pitch2(const float *xcorr, const float *y, int len, int max_pitch, int *best) {
    float Syy = 1;
    float best_num[2] = { -1, -1 }, best_den[2] = { 0, 0 };
    best[0] = 0; best[1] = 1;
    for (int j = 0; j < len; j++) Syy = Syy + y[j] * y[j];
    for (int i = 0; i < max_pitch; i++) {
        if (xcorr[i] > 0) {
            float x16 = xcorr[i];
            x16 *= 1e-12f;
            float num = x16 * x16;
            if (num * best_den[1] > best_num[1] * Syy) {
                if (num * best_den[0] > best_num[0] * Syy) {
                    best_num[1] = best_num[0]; best_den[1] = best_den[0]; best[1] = best[0];
                    best_num[0] = num; best_den[0] = Syy; best[0] = i;
                } else {
                    best_num[1] = num; best_den[1] = Syy; best[1] = i;
                }
            }
        }
        Syy += y[i + len] * y[i + len] - y[i] * y[i];
        Syy = Syy < 1 ? 1 : Syy;                  /* MAX32(1,Syy) */
    }
}

/* pitch_search (pitch.c:283-385) for len=960, max_pitch=588 */
static int pitch_search_port(const float *x_lp, const float *y, int len, int max_pitch) {
    int lag = len + max_pitch;
    float x4[240], y4[387], xc[294];
    int best[2] = { 0, 0 };
    for (int j = 0; j < len >> 2; j++) x4[j] = x_lp[2 * j];
    for (int j = 0; j < lag >> 2; j++) y4[j] = y[2 * j];
    for (int i = 0; i < max_pitch >> 2; i++) xc[i] = inner(x4, y4 + i, len >> 2);
    best_pitch2(xc, y4, len >> 2, max_pitch >> 2, best);
    for (int i = 0; i < max_pitch >> 1; i++) {
        xc[i] = 0;
        if (abs(i - 2 * best[0]) > 2 && abs(i - 2 * best[1]) > 2) continue;
        float sum = inner(x_lp, y + i, len >> 1);
        xc[i] = sum < -1 ? -1 : sum;              /* MAX32(-1,sum) */
    }
    best_pitch2(xc, y, len >> 1, max_pitch >> 1, best);
    int offset = 0;
    if (best[0] > 0 && best[0] < (max_pitch >> 1) - 1) {
        float a = xc[best[0] - 1], b = xc[best[0]], c = xc[best[0] + 1];
        if ((c - a) > .7f * (b - a)) offset = 1;
        else if ((a - c) > .7f * (b - c)) offset = -1;
    }
    return 2 * best[0] - offset;
}

static float pitch_gain(float xy, float xx, float yy) {   /* compute_pitch_gain, pitch.c:416-419 */
    return (float)(xy / sqrt(1 + xx * yy));
}

/* remove_doubling (pitch.c:423-526) */
static float remove_doubling_port(const float *xbuf, int maxperiod, int minperiod, int N, int *T0_,
                                  int prev_period, float prev_gain) {
    static const int second_check[16] = { 0, 0, 3, 2, 3, 2, 5, 2, 3, 2, 3, 2, 5, 2, 3, 2 };
    int minperiod0 = minperiod;
    maxperiod /= 2; minperiod /= 2; *T0_ /= 2; prev_period /= 2; N /= 2;
    const float *x = xbuf + maxperiod;
    if (*T0_ >= maxperiod) *T0_ = maxperiod - 1;
    int T, T0; T = T0 = *T0_;
    float yy_lookup[385];
    float xx = 0, xy = 0;
    for (int i = 0; i < N; i++) { xx = xx + x[i] * x[i]; xy = xy + x[i] * x[i - T0]; }
    yy_lookup[0] = xx;
    float yy = xx;
    for (int i = 1; i <= maxperiod; i++) {
        yy = yy + x[-i] * x[-i] - x[N - i] * x[N - i];
        yy_lookup[i] = yy < 0 ? 0 : yy;
    }
    yy = yy_lookup[T0];
    float best_xy = xy, best_yy = yy;
    float g, g0; g = g0 = pitch_gain(xy, xx, yy);
    for (int k = 2; k <= 15; k++) {
        int T1 = (2 * T0 + k) / (2 * k), T1b;
        if (T1 < minperiod) break;
        if (k == 2) T1b = (T1 + T0 > maxperiod) ? T0 : T0 + T1;
        else T1b = (2 * second_check[k] * T0 + k) / (2 * k);
        float xy1 = 0, xy2 = 0;
        for (int i = 0; i < N; i++) { xy1 = xy1 + x[i] * x[i - T1]; xy2 = xy2 + x[i] * x[i - T1b]; }
        xy = .5f * (xy1 + xy2);
        yy = .5f * (yy_lookup[T1] + yy_lookup[T1b]);
        float g1 = pitch_gain(xy, xx, yy);
        float cont;
        if (abs(T1 - prev_period) <= 1) cont = prev_gain;
        else if (abs(T1 - prev_period) <= 2 && 5 * k * k < T0) cont = .5f * prev_gain;
        else cont = 0;
        float a = .7f * g0 - cont;
        float thresh = .3f > a ? .3f : a;
        if (T1 < 3 * minperiod) { a = .85f * g0 - cont; thresh = .4f > a ? .4f : a; }
        else if (T1 < 2 * minperiod) { a = .9f * g0 - cont; thresh = .5f > a ? .5f : a; }
        if (g1 > thresh) { best_xy = xy; best_yy = yy; T = T1; g = g1; }
    }
    best_xy = best_xy < 0 ? 0 : best_xy;
    float pg;
    if (best_yy <= best_xy) pg = 1.0f;
    else pg = best_xy / (best_yy + 1);
    float xc[3];
    for (int k = 0; k < 3; k++) xc[k] = inner(x, x - (T + k - 1), N);
    int offset;
    if ((xc[2] - xc[0]) > .7f * (xc[1] - xc[0])) offset = 1;
    else if ((xc[0] - xc[2]) > .7f * (xc[1] - xc[2])) offset = -1;
    else offset = 0;
    if (pg > g) pg = g;
    *T0_ = 2 * T + offset;
    if (*T0_ < minperiod0) *T0_ = minperiod0;
    return pg;
}

/* ---- the RNN (rnn.c) ------------------------------------------------------ */
static float tansig(float x) {                    /* tansig_approx, rnn.c:41-67 */
    float y, dy, sign = 1;
    if (!(x < 8)) return 1;
    if (!(x > -8)) return -1;
    if (x != x) return 0;
    if (x < 0) { x = -x; sign = -1; }
    int i = (int)floor(.5f + 25 * x);
    x -= .04f * i;
    y = g_model.tansig[i];
    dy = 1 - y * y;
    y = y + x * dy * (1 - y * x);
    return sign * y;
}
static float sigmoid(float x) {                   /* sigmoid_approx, rnn.c:69-72: double arithmetic */
    return (float)(.5 + .5 * tansig((float)(.5 * x)));
}
#define WSCALE (1.f / 256)
enum { ACT_TANH = 0, ACT_SIGMOID = 1, ACT_RELU = 2 };

static void dense(const signed char *bias, const signed char *w, int M, int N, int act, float *out, const float *in) {
    for (int i = 0; i < N; i++) {                 /* compute_dense, rnn.c:79-107 */
        float sum = bias[i];
        for (int j = 0; j < M; j++) sum += w[j * N + i] * in[j];
        out[i] = WSCALE * sum;
    }
    for (int i = 0; i < N; i++) out[i] = act == ACT_SIGMOID ? sigmoid(out[i]) : tansig(out[i]);
}

static void gru(const signed char *bias, const signed char *w, const signed char *rw, int M, int N, float *state, const float *in) {
    float z[128], r[128], h[128];                 /* compute_gru, rnn.c:109-156 (ReLU candidate) */
    int stride = 3 * N;
    for (int i = 0; i < N; i++) {
        float sum = bias[i];
        for (int j = 0; j < M; j++) sum += w[j * stride + i] * in[j];
        for (int j = 0; j < N; j++) sum += rw[j * stride + i] * state[j];
        z[i] = sigmoid(WSCALE * sum);
    }
    for (int i = 0; i < N; i++) {
        float sum = bias[N + i];
        for (int j = 0; j < M; j++) sum += w[N + j * stride + i] * in[j];
        for (int j = 0; j < N; j++) sum += rw[N + j * stride + i] * state[j];
        r[i] = sigmoid(WSCALE * sum);
    }
    for (int i = 0; i < N; i++) {
        float sum = bias[2 * N + i];
        for (int j = 0; j < M; j++) sum += w[2 * N + j * stride + i] * in[j];
        for (int j = 0; j < N; j++) sum += rw[2 * N + j * stride + i] * state[j] * r[j];
        sum = WSCALE * sum;
        sum = sum < 0 ? 0 : sum;
        h[i] = z[i] * state[i] + (1 - z[i]) * sum;
    }
    for (int i = 0; i < N; i++) state[i] = h[i];
}

struct leaf_rnn {                                  /* DenoiseState, denoise.c:86-98 */
    float analysis_mem[FRAME];
    float cepstral_mem[CEPS][NB];
    int   memid;
    float synthesis_mem[FRAME];
    float pitch_buf[PBUF];
    float last_gain;
    int   last_period;
    float mem_hp_x[2];
    float lastg[NB];
    float vad_state[24], noise_state[48], denoise_state[96];
};

static void rnn_step(leaf_rnn *st, float *gains, float *vad, const float *feat) {   /* compute_rnn, rnn.c:160-178 */
    float dense_out[24], noise_in[90], den_in[114];
    dense(g_model.in_b, g_model.in_w, 42, 24, ACT_TANH, dense_out, feat);
    gru(g_model.vg_b, g_model.vg_w, g_model.vg_r, 24, 24, st->vad_state, dense_out);
    dense(g_model.vo_b, g_model.vo_w, 24, 1, ACT_SIGMOID, vad, st->vad_state);
    for (int i = 0; i < 24; i++) noise_in[i] = dense_out[i];
    for (int i = 0; i < 24; i++) noise_in[24 + i] = st->vad_state[i];
    for (int i = 0; i < 42; i++) noise_in[48 + i] = feat[i];
    gru(g_model.ng_b, g_model.ng_w, g_model.ng_r, 90, 48, st->noise_state, noise_in);
    for (int i = 0; i < 24; i++) den_in[i] = st->vad_state[i];
    for (int i = 0; i < 48; i++) den_in[24 + i] = st->noise_state[i];
    for (int i = 0; i < 42; i++) den_in[72 + i] = feat[i];
    gru(g_model.dg_b, g_model.dg_w, g_model.dg_r, 114, 96, st->denoise_state, den_in);
    dense(g_model.do_b, g_model.do_w, 96, 22, ACT_SIGMOID, gains, st->denoise_state);
}

/* ---- frame features (compute_frame_features, denoise.c:310-399) ------------ */
static int frame_features(leaf_rnn *st, cpx *X, cpx *P, float *Ex, float *Ep, float *Exp, float *features,
                          const float *in, int *pitch_out, float *gain_out) {
    float x[WIN], p[WIN], pbuf[PBUF >> 1], tmp[NB], Ly[NB];
    /* frame_analysis, denoise.c:295-308 */
    memcpy(x, st->analysis_mem, sizeof(float) * FRAME);
    memcpy(x + FRAME, in, sizeof(float) * FRAME);
    memcpy(st->analysis_mem, in, sizeof(float) * FRAME);
    window960(x);
    forward960(X, x);
    band_accumulate(Ex, X, NULL);

    memmove(st->pitch_buf, st->pitch_buf + FRAME, sizeof(float) * (PBUF - FRAME));
    memcpy(st->pitch_buf + PBUF - FRAME, in, sizeof(float) * FRAME);
    pitch_down(st->pitch_buf, pbuf, PBUF);
    int pitch_index = pitch_search_port(pbuf + (PMAX >> 1), pbuf, 960, PMAX - 3 * PMIN);
    pitch_index = PMAX - pitch_index;
    float gain = remove_doubling_port(pbuf, PMAX, PMIN, 960, &pitch_index, st->last_period, st->last_gain);
    st->last_period = pitch_index;
    st->last_gain = gain;
    *pitch_out = pitch_index; *gain_out = gain;
    for (int i = 0; i < WIN; i++) p[i] = st->pitch_buf[PBUF - WIN - pitch_index + i];
    window960(p);
    forward960(P, p);
    band_accumulate(Ep, P, NULL);
    band_accumulate(Exp, X, P);
    for (int i = 0; i < NB; i++) Exp[i] = (float)(Exp[i] / sqrt(.001 + Ex[i] * Ep[i]));
    dct22(tmp, Exp);
    for (int i = 0; i < NDELTA; i++) features[NB + 2 * NDELTA + i] = tmp[i];
    features[NB + 2 * NDELTA] = (float)(features[NB + 2 * NDELTA] - 1.3);
    features[NB + 2 * NDELTA + 1] = (float)(features[NB + 2 * NDELTA + 1] - 0.9);
    features[NB + 3 * NDELTA] = (float)(.01 * (pitch_index - 300));
    float logMax = -2, follow = -2, E = 0;
    for (int i = 0; i < NB; i++) {
        Ly[i] = (float)log10(1e-2 + Ex[i]);
        double a = follow - 1.5;                   /* MAX16(logMax-7, MAX16(follow-1.5, Ly[i])) */
        double inner_max = a > Ly[i] ? a : (double)Ly[i];
        float lm7 = logMax - 7;
        Ly[i] = (float)(lm7 > inner_max ? (double)lm7 : inner_max);
        logMax = logMax > Ly[i] ? logMax : Ly[i];
        a = follow - 1.5;
        follow = (float)(a > Ly[i] ? a : (double)Ly[i]);
        E += Ex[i];
    }
    if (E < 0.04) {                                /* float < double literal */
        memset(features, 0, sizeof(float) * NFEAT);
        return 1;
    }
    dct22(features, Ly);
    features[0] -= 12;
    features[1] -= 4;
    float *c0 = st->cepstral_mem[st->memid];
    float *c1 = (st->memid < 1) ? st->cepstral_mem[CEPS + st->memid - 1] : st->cepstral_mem[st->memid - 1];
    float *c2 = (st->memid < 2) ? st->cepstral_mem[CEPS + st->memid - 2] : st->cepstral_mem[st->memid - 2];
    for (int i = 0; i < NB; i++) c0[i] = features[i];
    st->memid++;
    for (int i = 0; i < NDELTA; i++) {
        features[i] = c0[i] + c1[i] + c2[i];
        features[NB + i] = c0[i] - c2[i];
        features[NB + NDELTA + i] = c0[i] - 2 * c1[i] + c2[i];
    }
    if (st->memid == CEPS) st->memid = 0;
    float spec_variability = 0;
    for (int i = 0; i < CEPS; i++) {
        float mindist = 1e15f;
        for (int j = 0; j < CEPS; j++) {
            float dist = 0;
            for (int k = 0; k < NB; k++) {
                float t = st->cepstral_mem[i][k] - st->cepstral_mem[j][k];
                dist += t * t;
            }
            if (j != i) mindist = mindist < dist ? mindist : dist;
        }
        spec_variability += mindist;
    }
    features[NB + 3 * NDELTA + 1] = (float)(spec_variability / CEPS - 2.1);
    return 0;
}

/* pitch_filter (denoise.c:422-456) */
static void pitch_filter_port(cpx *X, const cpx *P, const float *Ex, const float *Ep, const float *Exp, const float *g) {
    float r[NB], rf[NFREQ] = { 0 }, newE[NB], norm[NB], normf[NFREQ] = { 0 };
    for (int i = 0; i < NB; i++) {
        if (Exp[i] > g[i]) r[i] = 1;
        else r[i] = (float)((Exp[i] * Exp[i]) * (1 - (g[i] * g[i])) / (.001 + (g[i] * g[i]) * (1 - (Exp[i] * Exp[i]))));
        float c = 0 > r[i] ? 0 : r[i];
        c = 1 < c ? 1 : c;
        r[i] = (float)sqrt(c);
        r[i] = (float)(r[i] * sqrt(Ex[i] / (1e-8 + Ep[i])));
    }
    band_interp(rf, r);
    for (int i = 0; i < NFREQ; i++) { X[i].r += rf[i] * P[i].r; X[i].i += rf[i] * P[i].i; }
    band_accumulate(newE, X, NULL);
    for (int i = 0; i < NB; i++) norm[i] = (float)sqrt(Ex[i] / (1e-8 + newE[i]));
    band_interp(normf, norm);
    for (int i = 0; i < NFREQ; i++) { X[i].r *= normf[i]; X[i].i *= normf[i]; }
}

leaf_rnn *leaf_rnn_create(void) {
    model_load(); tables_init();
    return (leaf_rnn *)calloc(1, sizeof(leaf_rnn));
}
void leaf_rnn_destroy(leaf_rnn *r) { free(r); }

/* rnnoise_process_frame (denoise.c:458-494) */
float leaf_rnn_process(leaf_rnn *st, float *out, const float *in, orc_rnnoise_trace *tr) {
    cpx X[NFREQ], P[WIN];
    float x[FRAME], Ex[NB], Ep[NB], Exp[NB], features[NFEAT], g[NB], gf[NFREQ] = { 1 };
    float vad_prob = 0;
    static const float a_hp[2] = { -1.99599f, 0.99600f }, b_hp[2] = { -2, 1 };
    for (int i = 0; i < FRAME; i++) {              /* biquad, denoise.c:410-420 */
        float xi = in[i];
        float yi = in[i] + st->mem_hp_x[0];
        st->mem_hp_x[0] = (float)(st->mem_hp_x[1] + (b_hp[0] * (double)xi - a_hp[0] * (double)yi));
        st->mem_hp_x[1] = (float)(b_hp[1] * (double)xi - a_hp[1] * (double)yi);
        x[i] = yi;
    }
    int pitch_index = 0; float pgain = 0;
    int silence = frame_features(st, X, P, Ex, Ep, Exp, features, x, &pitch_index, &pgain);
    if (tr) {
        memcpy(tr->hp, x, sizeof x);
        memcpy(tr->X, X, sizeof(float) * 2 * NFREQ);
        memcpy(tr->Ex, Ex, sizeof Ex); memcpy(tr->Ep, Ep, sizeof Ep); memcpy(tr->Exp, Exp, sizeof Exp);
        tr->pitch_index = pitch_index; tr->pitch_gain = pgain; tr->silence = silence;
        memcpy(tr->features, features, sizeof features);
        memset(tr->gains, 0, sizeof tr->gains);
    }
    if (!silence) {
        rnn_step(st, g, &vad_prob, features);
        if (tr) memcpy(tr->gains, g, sizeof g);
        pitch_filter_port(X, P, Ex, Ep, Exp, g);
        for (int i = 0; i < NB; i++) {
            float alpha = .6f;
            float a = alpha * st->lastg[i];
            g[i] = g[i] > a ? g[i] : a;
            st->lastg[i] = g[i];
        }
        band_interp(gf, g);
        for (int i = 0; i < NFREQ; i++) { X[i].r *= gf[i]; X[i].i *= gf[i]; }
    }
    /* frame_synthesis + inverse_transform (denoise.c:401-408, 227-245) */
    {
        cpx xi[WIN], y[WIN];
        float t[WIN];
        for (int i = 0; i < NFREQ; i++) xi[i] = X[i];
        for (int i = NFREQ; i < WIN; i++) { xi[i].r = xi[WIN - i].r; xi[i].i = -xi[WIN - i].i; }
        fft960(xi, y);
        t[0] = WIN * y[0].r;
        for (int i = 1; i < WIN; i++) t[i] = WIN * y[WIN - i].r;
        window960(t);
        for (int i = 0; i < FRAME; i++) out[i] = t[i] + st->synthesis_mem[i];
        memcpy(st->synthesis_mem, t + FRAME, sizeof(float) * FRAME);
    }
    if (tr) { tr->vad_prob = vad_prob; memcpy(tr->out, out, sizeof(float) * FRAME); }
    return vad_prob;
}

const char *leaf_kind(void) { return "port"; }
