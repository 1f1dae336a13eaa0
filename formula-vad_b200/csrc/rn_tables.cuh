// rn_tables.cuh -- host-side construction of RNNoise's tables and weight matrices.
// Shared by the CUDA library (uploaded once per device) and by the host emulation used in CPU tests.
#pragma once
#include "rn_core.cuh"
#include <string>
#include <vector>

namespace fvad {
namespace rn {

struct HostTables {
    std::vector<float> half_window, dct, frac, omf, tansig, bias;
    std::vector<cpx> tw;
    std::vector<short> rev;
    std::vector<unsigned short> weights;     // bf16 bit patterns, layer matrices W[j][col] (rn_core.cuh rnn_layer)
    std::vector<unsigned> weights2;          // the same matrices with rows paired: W[j][col] | W[j+1][col] << 16
    double dct_scale = 0;
    Tables view() const {
        Tables t;
        t.half_window = half_window.data(); t.dct = dct.data(); t.tw = tw.data(); t.rev = rev.data();
        t.frac = frac.data(); t.omf = omf.data(); t.tansig = tansig.data(); t.dct_scale = dct_scale;
        return t;
    }
};

// Loads formula-vad_b200/data/rnnoise_model.bin (tools/extract_rnnoise_model.py) and computes the
// tables with the host libm, as the reference's check_init / opus_fft_alloc_twiddles do
// (denoise.c:168-182, kiss_fft.c:404-419).  Returns FVAD_OK or an error (message via set_error).
int host_tables_build(HostTables *out, const char *blob_path_or_null);
std::string default_blob_path();

}  // namespace rn
}  // namespace fvad
