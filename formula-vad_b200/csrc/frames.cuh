// frames.cuh -- small per-frame kernels around the transforms: channel RMS ratio
// (VAD.preAnalyzeSegment), 480->2048 re-framing bookkeeping (VAD.fftBufferStep), copies.
#pragma once
#include "common.cuh"

namespace fvad {

// preAnalyzeSegment + rmsVolume (VAD.zig:253-272, audio_utils.zig:14-24): for every
// (stream, frame) the min/max RMS ratio over channels.  work[row][row_stride], frames of
// frame_len samples starting at sample 0.  out[stream*out_stride + out_offset + frame].
int rms_ratio_launch(const float *d_work, size_t row_stride, size_t n_streams, int n_channels, size_t n_frames,
                     int frame_len, float *d_out, size_t out_stride, size_t out_offset, cudaStream_t stream);

// fftBufferStep bookkeeping (VAD.zig:298-348): STFT frame j (global number g0+j) is filled from
// consecutive 480-sample frames; volume_ratio accumulates ratio*share in f32 in piece order, vad is
// the last piece's.  ratio480/vad480 are indexed [stream*stride480 + hist + (global480 - base480)].
int reframe_launch(const float *d_ratio480, const float *d_vad480, size_t stride480, long long base480, int hist,
                   size_t n_streams, unsigned long long first_stft_frame, size_t n_stft_frames, int fft_size,
                   float *d_ratio_out, float *d_vad_out, size_t out_stride, cudaStream_t stream);

// dst[row*dst_stride + dst_off + i] = src[row*src_stride + src_off + i], i < n   (rows x n floats)
int copy_rows_launch(float *dst, size_t dst_stride, size_t dst_off, const float *src, size_t src_stride,
                     size_t src_off, size_t n_rows, size_t n, cudaStream_t stream);
// device input with separate stream/channel strides -> work rows
int gather_rows_launch(float *dst, size_t dst_stride, size_t dst_off, const float *src, size_t stream_stride,
                       size_t channel_stride, size_t src_off, size_t n_streams, int n_channels, size_t n,
                       cudaStream_t stream);

}  // namespace fvad

namespace fvad {
// Recorder (src/AudioPipeline/Recorder.zig:52-146): the original samples [start, start+len) of one history row, gathered
// out of the PCM history ring (absolute sample i of row r at d_hist[r*cap + i % cap]) into a packed staging buffer.
struct ClipPiece {
    size_t row;                     // history row = stream * n_channels + channel
    unsigned long long start;       // absolute sample index
    size_t len, dst_off;            // floats
};
int gather_clips_launch(const ClipPiece *d_pieces, size_t n_pieces, size_t max_len, const float *d_hist, size_t cap, float *d_dst,
                        cudaStream_t stream);
// arr[s][0..hist) = arr[s][n_new .. n_new+hist) for every stream (keeps the last `hist` entries as history)
int shift_hist_launch(float *arr, size_t stride, int hist, size_t n_new, size_t n_streams, cudaStream_t stream);
}
