// rnnoise.cuh -- batched RNNoise (lib/rnnoise/src/denoise.c:458-494 rnnoise_process_frame and
// everything below it) for many independent DenoiseStates x many consecutive frames.
#pragma once
#include "common.cuh"

namespace fvad {

struct RnnoiseBatch;

// n_states independent DenoiseStates (Denoiser.init -> rnnoise_create(NULL), Denoiser.zig:22-31),
// each able to take up to max_frames consecutive 480-sample frames per call.
int rnnoise_batch_create(size_t n_states, size_t max_frames, int n_slots, int device, RnnoiseBatch **out);
void rnnoise_batch_destroy(RnnoiseBatch *r);
int rnnoise_batch_reset(RnnoiseBatch *r, cudaStream_t stream);

// Process n_tframes time frames of every state.  A state serves n_channels channels in turn, as the
// reference does with its single shared DenoiseState (VAD.zig:274-296): the state's frame sequence is
// (t0,c0),(t0,c1),...,(t1,c0),...  Input rows are normalised [-1,1] PCM (Denoiser.zig:45-66 scaling is
// applied inside): sample i of (state s, channel c) at d_in[(s*n_channels+c)*in_stride + i].
// Outputs: d_out rows likewise (normalised, denoised) at d_out[row*out_stride + out_off + i];
// d_vad_min[s*vad_stride + vad_off + t] = min over channels of the frame's VAD probability, start 1
// (VAD.zig:284-293); d_vad_all (optional) [(s*n_channels+c)*vad_stride + vad_off + t].
int rnnoise_batch_process(RnnoiseBatch *r, const float *d_in, size_t in_stride, int n_channels, size_t n_tframes,
                          float *d_out, size_t out_stride, size_t out_off, float *d_vad_min, float *d_vad_all,
                          size_t vad_stride, size_t vad_off, cudaStream_t stream);


// The same work as stages that batch.cu issues on CUDA streams of their own (wavefront over tiles; B = B1 then B2): arrays
// that cross a stage boundary exist once per slot, cross-tile carries stay inside their stage.
//   A: biquad (sample-sequential)           B: analysis / pitch / features (frame-parallel)
//   C: RNN (frame-sequential)               D: gains, synthesis, overlap-add, VAD minimum
int rnnoise_n_slots(const RnnoiseBatch *r);
int rnnoise_stage_a(RnnoiseBatch *r, int slot, const float *d_in, size_t in_stride, int n_channels, size_t n_tframes, cudaStream_t st);
int rnnoise_stage_b(RnnoiseBatch *r, int slot, size_t n_vframes, cudaStream_t st);    // = b1 then b2 on one stream
int rnnoise_stage_b1(RnnoiseBatch *r, int slot, size_t n_vframes, cudaStream_t st);   // analysis + pitch correlations
int rnnoise_stage_b2(RnnoiseBatch *r, int slot, size_t n_vframes, cudaStream_t st);   // pitch decision scan + features
int rnnoise_stage_c(RnnoiseBatch *r, int slot, size_t n_vframes, cudaStream_t st);
int rnnoise_stage_d(RnnoiseBatch *r, int slot, int n_channels, size_t n_tframes, float *d_out, size_t out_stride, size_t out_off,
                    float *d_vad_min, float *d_vad_all, size_t vad_stride, size_t vad_off, cudaStream_t st);

}  // namespace fvad
