//! Drop-in for src/Denoiser.zig (init / denoise / getFrameSize / deinit, src/Denoiser.zig:22-70) over libfvad_cuda.so.
//! UNCOMPILED HERE (no Zig toolchain in this image).  One 480-sample frame per call is the worst way to drive a GPU --
//! the pipeline uses the batched form -- but the leaf API keeps its meaning: ONE DenoiseState per Denoiser, frames in
//! call order, bit-identical output and VAD probability.
const std = @import("std");
const Allocator = std.mem.Allocator;
const SplitSlice = @import("./structures/SplitSlice.zig").SplitSlice;
const c = @cImport(@cInclude("fvad.h"));

const Self = @This();

allocator: Allocator,
handle: ?*c.fvad_denoiser,

pub fn init(allocator: Allocator) !Self {
    var h: ?*c.fvad_denoiser = null;
    if (c.fvad_denoiser_create(&h) != c.FVAD_OK) return error.CudaError;
    return Self{ .allocator = allocator, .handle = h };
}

pub fn deinit(self: *Self) void {
    if (self.handle) |h| c.fvad_denoiser_destroy(h);
    self.handle = null;
}

/// Samples must be mono, 48 kHz, f32 in [-1, 1].  Returns the VAD probability; `result` gets the denoised frame.
pub fn denoise(self: *Self, samples: SplitSlice(f32), result: []f32) !f32 {
    const h = self.handle orelse return error.NotInitialized; // Denoiser.zig:46
    var vad: f32 = 0;
    const rc = c.fvad_denoiser_denoise(h, samples.first.ptr, samples.first.len, samples.second.ptr, samples.second.len,
        result.ptr, result.len, &vad);
    return switch (rc) {
        c.FVAD_OK => vad,
        c.FVAD_EINVALID_FRAME_SIZE => error.InvalidFrameSize, // :48-50
        c.FVAD_EINVALID_RESULT_SIZE => error.InvalidResultSize, // :52-54
        c.FVAD_ENOT_INITIALIZED => error.NotInitialized,
        else => error.CudaError,
    };
}

pub fn getFrameSize() usize { // :68-70
    return c.fvad_denoiser_frame_size();
}
