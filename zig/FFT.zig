//! Drop-in for src/FFT.zig (init / fft / binCount / binWidth / nyquistFreq / freqToBin / binToFreq / deinit,
//! src/FFT.zig:28-144) over libfvad_cuda.so.  UNCOMPILED HERE (no Zig toolchain in this image).
//! One transform per call crosses PCIe twice; the batched form (fvad_stft_batch over device memory) is what the pipeline
//! uses -- this file exists so that the leaf API keeps working for its other callers (Spectrogram.compute).
const std = @import("std");
const Allocator = std.mem.Allocator;
const SplitSlice = @import("./structures/SplitSlice.zig").SplitSlice;
const c = @cImport(@cInclude("fvad.h"));

const Self = @This();

allocator: Allocator,
n_fft: usize,
sample_rate: usize,
handle: *c.fvad_fft,

fn check(rc: c_int) !void {
    return switch (rc) {
        c.FVAD_OK => {},
        c.FVAD_EINVALID_FFT_SIZE => error.InvalidFFTSize, // FFT.zig:29-31
        c.FVAD_EINVALID_SAMPLES_LEN => error.InvalidSamplesLength, // :71-73
        c.FVAD_EINVALID_WINDOW_LEN => error.InvalidWindowLength, // :75-77
        c.FVAD_EINVALID_RESULT_LEN => error.InvalidResultLength, // :79-81
        c.FVAD_EOUT_OF_RANGE => error.OutOfRange, // :121-123,135-137
        c.FVAD_ENEGATIVE_FREQUENCY => error.NegativeFrequency, // :125-127
        c.FVAD_EOUT_OF_MEMORY => error.OutOfMemory,
        c.FVAD_EUNSUPPORTED => error.Unsupported, // an even n_fft this build has no kernel for (see fvad.h)
        else => error.CudaError,
    };
}

pub fn init(allocator: Allocator, n_fft: usize, sample_rate: usize) !Self {
    var h: ?*c.fvad_fft = null;
    try check(c.fvad_fft_create(n_fft, sample_rate, &h));
    return Self{ .allocator = allocator, .n_fft = n_fft, .sample_rate = sample_rate, .handle = h.? };
}

pub fn deinit(self: *Self) void {
    c.fvad_fft_destroy(self.handle);
}

/// FFT.fft (src/FFT.zig:70-98): windowed real transform of n_fft samples, normalised magnitudes of n_fft/2+1 bins.
pub fn fft(self: *Self, samples: SplitSlice(f32), window: []const f32, result: []f32) !void {
    try check(c.fvad_fft_run(self.handle, samples.first.ptr, samples.first.len, samples.second.ptr, samples.second.len,
        window.ptr, window.len, result.ptr, result.len));
}

pub fn binCount(self: Self) usize {
    return c.fvad_fft_bin_count(self.handle);
}
pub fn binWidth(self: Self) f32 {
    return c.fvad_fft_bin_width(self.handle);
}
pub fn nyquistFreq(self: Self) f32 {
    return c.fvad_fft_nyquist_freq(self.handle);
}
pub fn freqToBin(self: Self, freq: f32) !usize {
    var bin: usize = 0;
    try check(c.fvad_fft_freq_to_bin(self.handle, freq, &bin));
    return bin;
}
pub fn binToFreq(self: Self, bin: usize) !f32 {
    var f: f32 = 0;
    try check(c.fvad_fft_bin_to_freq(self.handle, bin, &f));
    return f;
}
