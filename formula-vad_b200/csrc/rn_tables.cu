// rn_tables.cu -- see rn_tables.cuh
#include "rn_tables.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>

namespace fvad {
namespace rn {

std::string default_blob_path() {
    const char *env = debug_knob("model_blob");
    if (env && *env) return env;
    Dl_info info;
    if (dladdr((void *)&default_blob_path, &info) && info.dli_fname) {
        std::string p = info.dli_fname;                 // <pkg>/libfvad_cuda.so or <pkg>/librn_emu.so
        size_t k = p.rfind('/');
        if (k != std::string::npos) return p.substr(0, k) + "/data/rnnoise_model.bin";
    }
    return "formula-vad_b200/data/rnnoise_model.bin";
}

namespace {
struct Blob {
    std::vector<unsigned char> raw;
    const void *find(const char *name, unsigned want, unsigned dtype) const {
        const unsigned n = *(const unsigned *)(raw.data() + 8);
        size_t off = 12 + (size_t)n * 40;
        for (unsigned a = 0; a < n; a++) {
            const unsigned char *e = raw.data() + 12 + (size_t)a * 40;
            const unsigned dt = *(const unsigned *)(e + 32), cnt = *(const unsigned *)(e + 36);
            const size_t bytes = (size_t)cnt * (dt ? 4 : 1);
            if (strncmp((const char *)e, name, 32) == 0) return (cnt == want && dt == dtype) ? raw.data() + off : nullptr;
            off += (bytes + 15) & ~(size_t)15;
        }
        return nullptr;
    }
};
unsigned short bf16_of_int8(signed char w) {
    float f = (float)w;
    unsigned u; memcpy(&u, &f, 4);
    return (unsigned short)(u >> 16);                   // exact: |w| <= 128 needs 8 significant bits
}
}  // namespace

int host_tables_build(HostTables *out, const char *blob_path) {
    const std::string path = blob_path ? std::string(blob_path) : default_blob_path();
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) return set_error(FVAD_ENOT_INITIALIZED, "cannot open RNNoise model blob %s", path.c_str());
    Blob b;
    fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    b.raw.resize((size_t)sz);
    const bool ok = fread(b.raw.data(), 1, (size_t)sz, fp) == (size_t)sz && sz > 12 && !memcmp(b.raw.data(), "FVADRNN1", 8);
    fclose(fp);
    if (!ok) return set_error(FVAD_ENOT_INITIALIZED, "bad RNNoise model blob %s", path.c_str());

    struct Src { const char *w, *r, *b; int M, N, gates; };
    const Src layers[6] = {
        {"input_dense_weights", nullptr, "input_dense_bias", 42, 24, 1},
        {"vad_gru_weights", "vad_gru_recurrent_weights", "vad_gru_bias", 24, 24, 3},
        {"vad_output_weights", nullptr, "vad_output_bias", 24, 1, 1},
        {"noise_gru_weights", "noise_gru_recurrent_weights", "noise_gru_bias", 90, 48, 3},
        {"denoise_gru_weights", "denoise_gru_recurrent_weights", "denoise_gru_bias", 114, 96, 3},
        {"denoise_output_weights", nullptr, "denoise_output_bias", 96, 22, 1},
    };
    out->weights.assign(kRnnWeights, 0);
    out->bias.assign(kRnnBiases, 0.f);
    for (int l = 0; l < 6; l++) {
        const RnnLayer L = rnn_layer(l);
        const Src &s = layers[l];
        const int cols = s.N * s.gates;
        if (cols != L.cols || L.M != s.M) return set_error(FVAD_ENOT_INITIALIZED, "layer table mismatch");
        const signed char *w = (const signed char *)b.find(s.w, (unsigned)(s.M * cols), 0);
        const signed char *bi = (const signed char *)b.find(s.b, (unsigned)cols, 0);
        const signed char *r = s.r ? (const signed char *)b.find(s.r, (unsigned)(s.N * cols), 0) : nullptr;
        if (!w || !bi || (s.r && !r)) return set_error(FVAD_ENOT_INITIALIZED, "model blob lacks layer %d arrays", l);
        for (int j = 0; j < s.M; j++)
            for (int c = 0; c < cols; c++) out->weights[L.w_off + j * cols + c] = bf16_of_int8(w[j * cols + c]);
        if (r)
            for (int j = 0; j < s.N; j++)
                for (int c = 0; c < cols; c++) out->weights[L.w_off + (s.M + j) * cols + c] = bf16_of_int8(r[j * cols + c]);
        for (int c = 0; c < cols; c++) out->bias[L.b_off + c] = (float)bi[c];
    }
    out->weights2.assign(kRnnWeights / 2, 0u);
    for (int l = 0; l < 6; l++) {
        const RnnLayer L = rnn_layer(l);
        if ((L.J & 1) || (L.M & 1) || (L.w_off & 1)) return set_error(FVAD_ENOT_INITIALIZED, "layer %d cannot be row-paired", l);
        for (int j = 0; j < L.J; j += 2)
            for (int c = 0; c < L.cols; c++)
                out->weights2[L.w_off / 2 + (j / 2) * L.cols + c] =
                    (unsigned)out->weights[L.w_off + j * L.cols + c] | ((unsigned)out->weights[L.w_off + (j + 1) * L.cols + c] << 16);
    }
    const float *ts = (const float *)b.find("tansig_table", 201, 1);
    if (!ts) return set_error(FVAD_ENOT_INITIALIZED, "model blob lacks tansig_table");
    out->tansig.assign(ts, ts + 201);

    out->half_window.resize(kFrame);
    for (int i = 0; i < kFrame; i++)                    // denoise.c:171-172
        out->half_window[i] = (float)sin(.5 * M_PI * sin(.5 * M_PI * (i + .5) / kFrame) * sin(.5 * M_PI * (i + .5) / kFrame));
    out->dct.resize(kNB * kNB);
    for (int i = 0; i < kNB; i++)                       // denoise.c:173-179
        for (int j = 0; j < kNB; j++) {
            out->dct[i * kNB + j] = (float)cos((i + .5) * j * M_PI / kNB);
            if (j == 0) out->dct[i * kNB + j] = (float)(out->dct[i * kNB + j] * sqrt(.5));
        }
    out->dct_scale = sqrt(2. / 22);
    out->tw.resize(kWin);
    for (int i = 0; i < kWin; i++) {                    // kiss_fft.c:404-419
        const double pi = 3.14159265358979323846264338327;
        const double phase = (-2 * pi / kWin) * i;
        out->tw[i].r = (float)cos(phase);
        out->tw[i].i = (float)sin(phase);
    }
    out->rev.resize(kWin);
    for (int n = 0; n < kWin; n++) {                    // kiss_fft.c:314-343 for radices 5,3,4,4,4
        const int n0 = n % 5, n1 = (n / 5) % 3, n2 = (n / 15) % 4, n3 = (n / 60) % 4, n4 = n / 240;
        out->rev[n] = (short)(192 * n0 + 64 * n1 + 16 * n2 + 4 * n3 + n4);
    }
    out->frac.assign(400, 0.f); out->omf.assign(400, 0.f);
    for (int i = 0; i < kNB - 1; i++) {                 // denoise.c:108-112
        const int band_size = (eband(i + 1) - eband(i)) << 2;
        for (int j = 0; j < band_size; j++) {
            const float frac = (float)j / band_size;
            out->frac[(eband(i) << 2) + j] = frac;
            out->omf[(eband(i) << 2) + j] = 1 - frac;
        }
    }
    return FVAD_OK;
}

}  // namespace rn
}  // namespace fvad
