// frames.cu -- see frames.cuh.
#include "frames.cuh"

namespace fvad {

__global__ void k_rms_ratio(const float *__restrict__ work, size_t row_stride, size_t n_streams, int C,
                            size_t n_frames, int frame_len, float *__restrict__ out, size_t out_stride,
                            size_t out_offset) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_streams * n_frames) return;
    // consecutive threads take consecutive frames of one stream
    const size_t s = id / n_frames, f = id - s * n_frames;
    float vol_min = 1.0f, vol_max = 0.0f;                       // VAD.zig:257-258
    for (int c = 0; c < C; c++) {
        const float *p = work + (s * C + c) * row_stride + f * (size_t)frame_len;
        float sum = 0.0f;                                       // in-order f32 sum of squares
        if ((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (frame_len & 3) == 0) {
            const float4 *p4 = reinterpret_cast<const float4 *>(p);
            for (int i = 0; i < frame_len / 4; i++) {
                float4 v = p4[i];
                sum += v.x * v.x; sum += v.y * v.y; sum += v.z * v.z; sum += v.w * v.w;
            }
        } else {
            for (int i = 0; i < frame_len; i++) sum += p[i] * p[i];
        }
        const float mean = sum / (float)frame_len;
        const float vol = sqrtf(mean);
        if (vol < vol_min) vol_min = vol;
        if (vol > vol_max) vol_max = vol;
    }
    out[s * out_stride + out_offset + f] = (vol_max == 0.0f) ? 0.0f : vol_min / vol_max;   // VAD.zig:267
}

int rms_ratio_launch(const float *d_work, size_t row_stride, size_t n_streams, int n_channels, size_t n_frames,
                     int frame_len, float *d_out, size_t out_stride, size_t out_offset, cudaStream_t stream) {
    const size_t n = n_streams * n_frames;
    if (!n) return FVAD_OK;
    FVAD_LAUNCH(k_rms_ratio, (unsigned)ceil_div(n, 64), 64, 0, stream, d_work, row_stride, n_streams, n_channels,
                n_frames, frame_len, d_out, out_stride, out_offset);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

__global__ void k_reframe(const float *__restrict__ ratio480, const float *__restrict__ vad480, size_t stride480,
                          long long base480, int hist, size_t n_streams, unsigned long long first_stft,
                          size_t n_stft, int fft_size, float *__restrict__ ratio_out, float *__restrict__ vad_out,
                          size_t out_stride) {
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_streams * n_stft) return;
    const size_t s = id / n_stft, j = id - s * n_stft;
    const unsigned long long begin = (first_stft + j) * (unsigned long long)fft_size, end = begin + fft_size;
    float acc = 0.0f, vad = 0.0f;
    unsigned long long pos = begin;
    while (pos < end) {
        const unsigned long long q = pos / 480;                 // global 480-frame number
        const unsigned long long qend = (q + 1) * 480;
        const unsigned long long upto = qend < end ? qend : end;
        const float share = (float)(upto - pos) / (float)fft_size;      // VAD.zig:320
        const size_t at = s * stride480 + (size_t)(hist + ((long long)q - base480));
        acc += ratio480[at] * share;                            // VAD.zig:326
        vad = vad480[at];                                       // VAD.zig:330: the last piece's value
        pos = upto;
    }
    ratio_out[s * out_stride + j] = acc;
    vad_out[s * out_stride + j] = vad;
}

int reframe_launch(const float *d_ratio480, const float *d_vad480, size_t stride480, long long base480, int hist,
                   size_t n_streams, unsigned long long first_stft_frame, size_t n_stft_frames, int fft_size,
                   float *d_ratio_out, float *d_vad_out, size_t out_stride, cudaStream_t stream) {
    const size_t n = n_streams * n_stft_frames;
    if (!n) return FVAD_OK;
    FVAD_LAUNCH(k_reframe, (unsigned)ceil_div(n, 128), 128, 0, stream, d_ratio480, d_vad480, stride480, base480, hist,
                n_streams, first_stft_frame, n_stft_frames, fft_size, d_ratio_out, d_vad_out, out_stride);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

__global__ void k_copy_rows(float *__restrict__ dst, size_t dst_stride, size_t dst_off, const float *__restrict__ src,
                            size_t src_stride, size_t src_off, size_t n) {
    const size_t row = blockIdx.y;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[row * dst_stride + dst_off + i] = src[row * src_stride + src_off + i];
}

int copy_rows_launch(float *dst, size_t dst_stride, size_t dst_off, const float *src, size_t src_stride,
                     size_t src_off, size_t n_rows, size_t n, cudaStream_t stream) {
    if (!n_rows || !n) return FVAD_OK;
    unsigned gx = (unsigned)ceil_div(n, 256 * 8);
    if (gx > 1024) gx = 1024;
    if (gx < 1) gx = 1;
    for (size_t r0 = 0; r0 < n_rows; r0 += 65535) {
        size_t nr = n_rows - r0 < 65535 ? n_rows - r0 : 65535;
        FVAD_LAUNCH(k_copy_rows, dim3(gx, (unsigned)nr), 256, 0, stream, dst + r0 * dst_stride, dst_stride, dst_off,
                    src + r0 * src_stride, src_stride, src_off, n);
    }
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

__global__ void k_gather_rows(float *__restrict__ dst, size_t dst_stride, size_t dst_off, const float *__restrict__ src,
                              size_t stream_stride, size_t channel_stride, size_t src_off, int C, size_t n) {
    const size_t row = blockIdx.y;
    const size_t s = row / C, c = row - s * C;
    const float *p = src + s * stream_stride + c * channel_stride + src_off;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[row * dst_stride + dst_off + i] = p[i];
}

int gather_rows_launch(float *dst, size_t dst_stride, size_t dst_off, const float *src, size_t stream_stride,
                       size_t channel_stride, size_t src_off, size_t n_streams, int n_channels, size_t n,
                       cudaStream_t stream) {
    const size_t n_rows = n_streams * n_channels;
    if (!n_rows || !n) return FVAD_OK;
    if (n_rows > 65535) return set_error(FVAD_EUNSUPPORTED, "more than 65535 rows per gather");
    unsigned gx = (unsigned)ceil_div(n, 256 * 8);
    if (gx > 1024) gx = 1024;
    FVAD_LAUNCH(k_gather_rows, dim3(gx, (unsigned)n_rows), 256, 0, stream, dst, dst_stride, dst_off, src, stream_stride,
                channel_stride, src_off, n_channels, n);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

}  // namespace fvad

namespace fvad {
// Recorder clips: piece i = `len` samples of history row `row` from absolute sample `start` (ring position start % cap)
// to dst + dst_off.  One CTA column per piece.
__global__ void k_gather_clips(const ClipPiece *__restrict__ pieces, const float *__restrict__ hist, size_t cap,
                               float *__restrict__ dst) {
    const ClipPiece p = pieces[blockIdx.y];
    const float *row = hist + p.row * cap;
    size_t pos = (size_t)((p.start + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) % cap);
    const size_t step = (size_t)gridDim.x * blockDim.x, wrap = step % cap;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.len; i += step) {
        dst[p.dst_off + i] = row[pos];
        pos += wrap;
        if (pos >= cap) pos -= cap;
    }
}

int gather_clips_launch(const ClipPiece *d_pieces, size_t n_pieces, size_t max_len, const float *d_hist, size_t cap, float *d_dst,
                        cudaStream_t stream) {
    if (!n_pieces || !max_len) return FVAD_OK;
    unsigned gx = (unsigned)ceil_div(max_len, 256 * 8);
    if (gx > 256) gx = 256;
    for (size_t p0 = 0; p0 < n_pieces; p0 += 65535) {
        const size_t np = n_pieces - p0 < 65535 ? n_pieces - p0 : 65535;
        FVAD_LAUNCH(k_gather_clips, dim3(gx, (unsigned)np), 256, 0, stream, d_pieces + p0, d_hist, cap, d_dst);
    }
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}

__global__ void k_shift_hist(float *arr, size_t stride, int hist, size_t n_new, size_t n_streams) {
    const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    float tmp[16];
    for (int i = 0; i < hist; i++) tmp[i] = arr[s * stride + n_new + i];
    for (int i = 0; i < hist; i++) arr[s * stride + i] = tmp[i];
}
int shift_hist_launch(float *arr, size_t stride, int hist, size_t n_new, size_t n_streams, cudaStream_t stream) {
    if (hist > 16) return set_error(FVAD_EINVALID_ARGUMENT, "hist > 16");
    FVAD_LAUNCH(k_shift_hist, (unsigned)ceil_div(n_streams, 128), 128, 0, stream, arr, stride, hist, n_new, n_streams);
    FVAD_CUDA_TRY(cudaGetLastError());
    return FVAD_OK;
}
}  // namespace fvad
