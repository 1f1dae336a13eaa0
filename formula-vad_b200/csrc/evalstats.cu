// evalstats.cu -- Evaluator statistics for many (stream, config) result sets at once.
// Reference: src/Evaluator.zig:90-156 (matching), src/Evaluator/statistics.zig:85-113,181-284 (TP/FP/FN seconds with
// extrusion and gap filling), src/Evaluator/SpeechSegment.zig:22-49.  One thread per result set; every sum is the
// reference's f32 accumulation in the reference's order.
#include "common.cuh"
#include <vector>

namespace fvad {
namespace {

__device__ __forceinline__ float overlap_with(float af, float at, float bf, float bt) {      // SpeechSegment.zig:22-27
    return fminf(at, bt) - fmaxf(af, bf);
}
__device__ __forceinline__ float seg_sec(uint64_t sample, float sr) { return (float)sample / sr; }   // SimulationInstance.zig:226-227

__global__ void k_eval_stats(const fvad_segment *__restrict__ segs, size_t n_sets, size_t max_segs,
                             const uint32_t *__restrict__ counts, const uint32_t *__restrict__ set_ref,
                             const float *__restrict__ ref_from, const float *__restrict__ ref_to,
                             const uint32_t *__restrict__ ref_offsets, float sample_rate, fvad_stat_config cfg,
                             fvad_single_stats *__restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sets) return;
    const fvad_segment *in = segs + i * max_segs;
    const uint32_t n_in = counts[i] < max_segs ? counts[i] : (uint32_t)max_segs;
    const uint32_t r0 = ref_offsets[set_ref[i]], r1 = ref_offsets[set_ref[i] + 1];
    fvad_single_stats st;
    st.total_positives_sec = st.true_positives_sec = st.false_positives_sec = st.false_negatives_sec = 0.f;

    for (uint32_t a = 0; a < n_in; a++) {                               // statistics.zig:88-94
        const float sf = seg_sec(in[a].sample_from, sample_rate), stt = seg_sec(in[a].sample_to, sample_rate);
        uint32_t m = 0;                                                 // findOverlapping: overlap > 0, in reference order
        for (uint32_t r = r0; r < r1; r++) m += overlap_with(sf, stt, ref_from[r], ref_to[r]) > 0.0f;
        // extrudeSegments (statistics.zig:225-250) + calcOverlapMany (:279-283), streamed over the matches
        float ov = 0.f, pf = 0.f, pt = 0.f;
        uint32_t k = 0;
        bool have = false;
        for (uint32_t r = r0; r < r1 && k < m; r++) {
            if (!(overlap_with(sf, stt, ref_from[r], ref_to[r]) > 0.0f)) continue;
            float nf = ref_from[r], nt = ref_to[r];
            if (k == 0) nf -= cfg.extrude_start;                        // first.from_sec -= extrude_start
            if (k == m - 1) nt += cfg.extrude_end;                      // last.to_sec += extrude_end
            if (have) {
                if (nf - pt <= cfg.fill_gaps) pt = nf;                  // bridge the gap to the next match
                ov += fmaxf(0.0f, overlap_with(sf, stt, pf, pt));
            }
            pf = nf; pt = nt; have = true; k++;
        }
        if (have) ov += fmaxf(0.0f, overlap_with(sf, stt, pf, pt));
        const float dur = stt - sf;
        const float fp = dur - ov;                                      // calcFalsePositiveSec, :193-205
        const float tp = dur - fp;                                      // calcTruePositiveSec, :207-216
        st.false_positives_sec += fp;
        st.true_positives_sec += tp;
        st.total_positives_sec += tp;
    }
    for (uint32_t r = r0; r < r1; r++) {                                // statistics.zig:96-102
        const float rf = ref_from[r], rt = ref_to[r];
        const float dur = rt - rf;
        if (dur < cfg.ignore_shorter_than_sec) continue;
        float ov = 0.f;                                                 // calcOverlapWithMatches, :273-277
        for (uint32_t a = 0; a < n_in; a++) {
            const float o = overlap_with(rf, rt, seg_sec(in[a].sample_from, sample_rate), seg_sec(in[a].sample_to, sample_rate));
            if (o > 0.0f) ov += fmaxf(0.0f, o);
        }
        const float fn = dur - ov;
        st.false_negatives_sec += fn;
        st.total_positives_sec += fn;
    }
    st.true_positive_rate = st.true_positives_sec / st.total_positives_sec;                       // :104-107
    st.false_negative_rate = st.false_negatives_sec / st.total_positives_sec;
    st.false_discovery_rate = st.false_positives_sec / (st.false_positives_sec + st.true_positives_sec);
    st.precision = st.true_positives_sec / (st.true_positives_sec + st.false_positives_sec);
    const float beta = 0.7f;                                                                      // :109-111
    st.f_score_beta = beta;
    st.f_score = (1 + beta * beta) * (st.precision * st.true_positive_rate) / (beta * beta * st.precision + st.true_positive_rate);
    st.fm_index = sqrtf(st.precision * st.true_positive_rate);
    out[i] = st;
}

template <class T> int to_device(T **d, const T *h, size_t n) {
    *d = nullptr;
    if (n == 0) n = 1;
    if (cudaMalloc(d, sizeof(T) * n) != cudaSuccess)
        return set_error(FVAD_EOUT_OF_MEMORY, "device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (h && cudaMemcpy(*d, h, sizeof(T) * n, cudaMemcpyHostToDevice) != cudaSuccess)
        return set_error(FVAD_ECUDA, "H2D: %s", cudaGetErrorString(cudaGetLastError()));
    return FVAD_OK;
}

}  // namespace
}  // namespace fvad

using namespace fvad;

static int eval_stats_run(const fvad_segment *segs, bool on_device, size_t n_sets, size_t max_segs, const uint32_t *counts,
                          const uint32_t *set_ref, const float *ref_from, const float *ref_to, const uint32_t *ref_offsets,
                          size_t n_refs, size_t sample_rate, const fvad_stat_config *config, fvad_single_stats *out) {
    if (!segs || !counts || !set_ref || !ref_offsets || !config || !out || !n_sets || !n_refs || !max_segs)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_eval_stats_batch: null/zero argument");
    for (size_t i = 0; i < n_sets; i++)
        if (set_ref[i] >= n_refs) return set_error(FVAD_EOUT_OF_RANGE, "set %zu names reference list %u of %zu", i, set_ref[i], n_refs);
    const size_t n_lab = ref_offsets[n_refs];
    if (n_lab && (!ref_from || !ref_to)) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_eval_stats_batch: null reference labels");
    fvad_segment *d_segs = nullptr; uint32_t *d_counts = nullptr, *d_set_ref = nullptr, *d_off = nullptr;
    float *d_from = nullptr, *d_to = nullptr; fvad_single_stats *d_out = nullptr;
    int rc = FVAD_OK;
    if (!on_device) {
        rc = to_device(&d_segs, segs, n_sets * max_segs);
        if (rc == FVAD_OK) rc = to_device(&d_counts, counts, n_sets);
    }
    if (rc == FVAD_OK) rc = to_device(&d_set_ref, set_ref, n_sets);
    if (rc == FVAD_OK) rc = to_device(&d_off, ref_offsets, n_refs + 1);
    if (rc == FVAD_OK) rc = to_device(&d_from, ref_from, n_lab);
    if (rc == FVAD_OK) rc = to_device(&d_to, ref_to, n_lab);
    if (rc == FVAD_OK) rc = to_device<fvad_single_stats>(&d_out, nullptr, n_sets);
    if (rc == FVAD_OK) {
        FVAD_LAUNCH(k_eval_stats, (unsigned)ceil_div(n_sets, 128), 128, 0, 0, on_device ? segs : d_segs, n_sets, max_segs,
                    on_device ? counts : d_counts, d_set_ref, d_from, d_to, d_off, (float)sample_rate, *config, d_out);
        if (cudaMemcpy(out, d_out, sizeof(fvad_single_stats) * n_sets, cudaMemcpyDeviceToHost) != cudaSuccess)
            rc = set_error(FVAD_ECUDA, "k_eval_stats: %s", cudaGetErrorString(cudaGetLastError()));
    }
    cudaFree(d_segs); cudaFree(d_counts); cudaFree(d_set_ref); cudaFree(d_off); cudaFree(d_from); cudaFree(d_to); cudaFree(d_out);
    return rc;
}

extern "C" int fvad_eval_stats_batch(const fvad_segment *segs, size_t n_sets, size_t max_segs, const uint32_t *counts,
                                     const uint32_t *set_ref, const float *ref_from, const float *ref_to,
                                     const uint32_t *ref_offsets, size_t n_refs, size_t sample_rate,
                                     const fvad_stat_config *config, int device, fvad_single_stats *out) {
    if (device < 0) FVAD_CUDA_TRY(cudaGetDevice(&device));
    int prev = 0;
    FVAD_CUDA_TRY(cudaGetDevice(&prev));
    FVAD_CUDA_TRY(cudaSetDevice(device));
    const int rc = eval_stats_run(segs, false, n_sets, max_segs, counts, set_ref, ref_from, ref_to, ref_offsets, n_refs, sample_rate, config, out);
    cudaSetDevice(prev);
    return rc;
}

extern "C" int fvad_eval_stats_batch_device(const fvad_segment *d_segs, size_t n_sets, size_t max_segs, const uint32_t *d_counts,
                                            const uint32_t *set_ref, const float *ref_from, const float *ref_to,
                                            const uint32_t *ref_offsets, size_t n_refs, size_t sample_rate,
                                            const fvad_stat_config *config, fvad_single_stats *out) {
    return eval_stats_run(d_segs, true, n_sets, max_segs, d_counts, set_ref, ref_from, ref_to, ref_offsets, n_refs, sample_rate, config, out);
}
