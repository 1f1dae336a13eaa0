// vadm_sweep.cu -- many VADMachine configurations over SHARED features (VAD.stateMachineStep's alt machines,
// src/AudioPipeline/VAD.zig:171-183,375-381; BASELINE config 5): one device thread per (stream, config) scans ALL records
// of a run in one launch, its state in registers (vadm_step.cuh).  Ring entry i of scan `id` lives at
// rings[i * n_scans + id]: the 32 scans of a warp touch one 256-byte line per ring index, and the 32 configs of a warp
// read the same record of the same stream (one broadcast load).  Each scan writes its own list, in emission order.
#include "vadm_step.cuh"
#include <algorithm>

namespace fvad {

void vadm_sweep_rows(const VadmParams *params, size_t n_configs, int *max_long, int *max_short, int *max_ratio) {
    int a = 0, b = 0, c = 0;
    for (size_t k = 0; k < n_configs; k++) {
        a = std::max(a, params[k].long_len); b = std::max(b, params[k].short_len); c = std::max(c, params[k].ratio_len);
    }
    *max_long = a; *max_short = b; *max_ratio = c;
}

__global__ void __launch_bounds__(32)
k_vadm_sweep(const VadmParams *__restrict__ params, double *__restrict__ rings, int max_long, int max_short,
             size_t n_streams, int n_configs, VadmSweepIo io) {
    const size_t n_scans = n_streams * (size_t)n_configs;
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_scans) return;
    const size_t s = id / n_configs;
    const VadmParams P = params[id - s * n_configs];
    const RingCol rl{rings + id, n_scans};
    const RingCol rs{rings + (size_t)max_long * n_scans + id, n_scans};
    const RingCol rr{rings + ((size_t)max_long + max_short) * n_scans + id, n_scans};
    VadmRegs S;
    vadm_regs_init(P, rl, rs, rr, S);        // the matrix keeps whatever the previous sweep left: every column used is rewritten
    const double k_rel = vadm_k_rel(P.long_len);
    const float *mv = io.min_volume + (size_t)P.band * io.band_stride + s * io.rec_stride;
    const float *rt = io.vol_ratio + s * io.rec_stride;
    const float *vd = io.vad ? io.vad + s * io.rec_stride : nullptr;
    fvad_segment *out = io.segs + id * io.max_segs;
    uint32_t n_out = 0;

    constexpr int kChunk = 8;
    for (size_t f0 = 0; f0 < io.n_records; f0 += kChunk) {
        float c_mv[kChunk], c_rt[kChunk], c_vd[kChunk];
#pragma unroll
        for (int j = 0; j < kChunk; j++) {                        // the chunk's records: independent loads, issued together
            const size_t f = f0 + j < io.n_records ? f0 + j : io.n_records - 1;
            c_mv[j] = __ldg(mv + f); c_rt[j] = __ldg(rt + f); c_vd[j] = vd ? __ldg(vd + f) : 0.0f;
        }
#pragma unroll 1
        for (int j = 0; j < kChunk && f0 + j < io.n_records; j++) {
            VadmStepOut o;
            vadm_step(P, k_rel, rl, rs, rr, S, c_mv[j], c_rt[j], c_vd[j], io.first_index + (f0 + j) * io.index_step, o);
            if (o.emit) {
                if (n_out < io.max_segs) out[n_out] = o.seg;
                n_out++;
            }
        }
    }
    io.counts[id] = n_out;
    if (io.fallbacks) io.fallbacks[id] = S.fallbacks;
}

int vadm_sweep_launch(const VadmParams *d_params, double *d_rings, int max_long, int max_short, size_t n_streams, int n_configs,
                      const VadmSweepIo &io, cudaStream_t stream) {
    const size_t n_scans = n_streams * (size_t)n_configs;
    if (!n_scans || !io.n_records) return FVAD_OK;
    FVAD_LAUNCH(k_vadm_sweep, (unsigned)ceil_div(n_scans, 32), 32, 0, stream, d_params, d_rings, max_long, max_short, n_streams,
                n_configs, io);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad
