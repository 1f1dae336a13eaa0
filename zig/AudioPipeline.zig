//! Drop-in for src/AudioPipeline.zig over libfvad_cuda.so (include/fvad.h).
//!
//! UNCOMPILED HERE: neither this image nor the GPU box has a Zig toolchain (the reference pins 0.11.0-dev.3198,
//! README.md:64).  Kept declaration for declaration with the reference so that a maintainer can drop it in:
//! `Config`, `Callbacks`, `init`, `deinit`, `pushSamples` keep their signatures (src/AudioPipeline.zig:15-120); the public
//! field path `pipeline.vad.vad_machine.vad_segments` (src/simulator/SimulationInstance.zig:211) becomes
//! `pipeline.vadSegments(0)`.  The Python/ctypes host (formula-vad_b200/fvad/__init__.py) drives the same C ABI call for
//! call and is what the parity tests exercise.
const std = @import("std");
const Allocator = std.mem.Allocator;
const c = @cImport(@cInclude("fvad.h"));
pub const VAD = @import("./AudioPipeline/VAD.zig"); // only Config / VADSpeechSegment / VADMachine.Config are used
pub const AudioBuffer = @import("./audio_utils/AudioBuffer.zig");

const Self = @This();

pub const Callbacks = struct { // unchanged: src/AudioPipeline.zig:15-18
    ctx: *anyopaque,
    on_recording: ?*const fn (ctx: *anyopaque, recording: *const AudioBuffer) void,
};

pub const Config = struct { // unchanged: src/AudioPipeline.zig:20-26
    sample_rate: usize,
    n_channels: usize,
    buffer_length: ?usize = null,
    vad_config: VAD.Config = .{},
    skip_processing: bool = false,
};

allocator: Allocator,
config: Config,
handle: *c.fvad_pipeline,
row_ptrs: []?[*]const f32,
callbacks: ?Callbacks = null,

pub const Error = error{
    InvalidSampleRate, InvalidFFTSize, ChannelCountMismatch, OutOfMemory, OutOfRange, NegativeFrequency,
    InvalidArgument, Unsupported, CudaError,
};

/// fvad_status -> the reference's error names (fvad_status_name gives the same strings).
pub fn check(rc: c_int) Error!void {
    return switch (rc) {
        c.FVAD_OK => {},
        c.FVAD_EINVALID_SAMPLE_RATE => error.InvalidSampleRate, // VAD.zig:101-104
        c.FVAD_EINVALID_FFT_SIZE => error.InvalidFFTSize, // FFT.zig:29-31
        c.FVAD_ECHANNEL_COUNT_MISMATCH => error.ChannelCountMismatch, // SegmentWriter.zig:63-65
        c.FVAD_EOUT_OF_MEMORY => error.OutOfMemory,
        c.FVAD_EOUT_OF_RANGE => error.OutOfRange,
        c.FVAD_ENEGATIVE_FREQUENCY => error.NegativeFrequency,
        c.FVAD_EINVALID_ARGUMENT => error.InvalidArgument,
        c.FVAD_EUNSUPPORTED => error.Unsupported,
        else => error.CudaError, // there is no CPU fallback
    };
}

pub fn machineConfigToC(m: VAD.VADMachine.Config) c.fvad_vad_machine_config {
    return .{
        .speech_min_freq = m.speech_min_freq,
        .speech_max_freq = m.speech_max_freq,
        .long_term_speech_avg_sec = m.long_term_speech_avg_sec,
        .has_initial_long_term_avg = @boolToInt(m.initial_long_term_avg != null),
        .initial_long_term_avg = m.initial_long_term_avg orelse 0,
        .short_term_speech_avg_sec = m.short_term_speech_avg_sec,
        .speech_threshold_factor = m.speech_threshold_factor,
        .channel_vol_ratio_avg_sec = m.channel_vol_ratio_avg_sec,
        .channel_vol_ratio_threshold = m.channel_vol_ratio_threshold,
        .min_consecutive_sec_to_open = m.min_consecutive_sec_to_open,
        .max_speech_gap_sec = m.max_speech_gap_sec,
        .min_vad_duration_sec = m.min_vad_duration_sec,
    };
}

/// The C side calls this from inside fvad_pipeline_push / fvad_batch_sync, on the calling thread, once per finished
/// recording -- where the reference calls `cb(cb_config.ctx, audio_buffer)` (src/AudioPipeline.zig:182-188).  The clip is
/// borrowed for the duration of the call, exactly as the reference frees its AudioBuffer right after the callback.
fn onRecordingTrampoline(ctx: ?*anyopaque, stream: usize, clip: [*c]const c.fvad_clip) callconv(.C) void {
    _ = stream; // a pipeline is a batch of one
    const self = @ptrCast(*Self, @alignCast(@alignOf(Self), ctx.?));
    const cbs = self.callbacks orelse return;
    const cb = cbs.on_recording orelse return;
    if (clip.*.lost != 0 or clip.*.pcm == null) return; // the history ring was too short: nothing to hand over
    var channel = [_][]f32{@intToPtr([*]f32, @ptrToInt(clip.*.pcm))[0..clip.*.length]};
    const audio_buffer = AudioBuffer{ // Recorder.segmentToAudioBuffer, src/AudioPipeline/Recorder.zig:121-146
        .allocator = self.allocator,
        .length = clip.*.length,
        .sample_rate = clip.*.sample_rate,
        .n_channels = 1,
        .duration_seconds = clip.*.duration_seconds,
        .channel_pcm_buf = &channel,
        .global_start_frame_number = clip.*.global_start_frame_number,
    };
    cb(cbs.ctx, &audio_buffer);
}

pub fn init(allocator: Allocator, config: Config, callbacks: ?Callbacks) !*Self {
    var cfg: c.fvad_config = undefined;
    c.fvad_default_config(&cfg, config.sample_rate, config.n_channels);
    cfg.buffer_length = config.buffer_length orelse 0;
    cfg.vad_config.fft_size = config.vad_config.fft_size;
    cfg.vad_config.use_denoiser = @boolToInt(config.vad_config.use_denoiser);
    cfg.vad_config.vad_machine_config = machineConfigToC(config.vad_config.vad_machine_config);
    const alt = config.vad_config.alt_vad_machine_configs orelse &[_]VAD.VADMachine.Config{};
    var alts = try allocator.alloc(c.fvad_vad_machine_config, alt.len);
    defer allocator.free(alts); // the library copies them
    for (alt, 0..) |m, i| alts[i] = machineConfigToC(m);
    cfg.vad_config.alt_vad_machine_configs = alts.ptr;
    cfg.vad_config.n_alt_vad_machine_configs = alts.len;
    cfg.skip_processing = @boolToInt(config.skip_processing);

    var self = try allocator.create(Self);
    errdefer allocator.destroy(self);
    var row_ptrs = try allocator.alloc(?[*]const f32, config.n_channels);
    errdefer allocator.free(row_ptrs);
    var h: ?*c.fvad_pipeline = null;
    try check(c.fvad_pipeline_create(&cfg, &h));
    errdefer c.fvad_pipeline_destroy(h);
    self.* = .{ .allocator = allocator, .config = config, .handle = h.?, .row_ptrs = row_ptrs, .callbacks = callbacks };
    if (callbacks) |cbs| {
        if (cbs.on_recording != null) {
            // history = 0: every pushed sample stays in HBM, so no clip is ever lost (a live caller that syncs often would
            // pass a bound, e.g. 120 s of samples)
            try check(c.fvad_batch_set_recording(self.handle, 0, onRecordingTrampoline, self));
        }
    }
    return self;
}

pub fn deinit(self: *Self) void {
    c.fvad_pipeline_destroy(self.handle);
    self.allocator.free(self.row_ptrs);
    self.allocator.destroy(self);
}

/// src/AudioPipeline.zig:86-120.  The samples are borrowed until the call returns; the stages run on the GPU behind the
/// caller's back and recordings / segments surface at `sync` (or inside a later push).
pub fn pushSamples(self: *Self, channel_pcm: []const []const f32) !u64 {
    if (channel_pcm.len != self.config.n_channels) return error.ChannelCountMismatch;
    for (channel_pcm, 0..) |ch, i| self.row_ptrs[i] = ch.ptr;
    var first: u64 = 0;
    try check(c.fvad_pipeline_push(self.handle, @ptrCast([*c]const [*c]const f32, self.row_ptrs.ptr), channel_pcm[0].len, &first));
    return first;
}

/// Wait for everything pushed so far: finished recordings are delivered to `on_recording` from inside this call.
pub fn sync(self: *Self) !void {
    try check(c.fvad_batch_sync(self.handle));
}

/// replaces `pipeline.vad.vad_machine.vad_segments` (machine_idx 0) and `pipeline.vad.alt_vad_machines.?[i - 1].vad_segments`
/// (src/simulator/SimulationInstance.zig:211; VAD.zig:94-95).  fvad_segment has VADSpeechSegment's layout (VAD.zig:25-30).
pub fn vadSegments(self: *Self, machine_idx: usize) ![]const VAD.VADSpeechSegment {
    var p: [*c]const c.fvad_segment = null;
    var n: usize = 0;
    try check(c.fvad_pipeline_segments(self.handle, @intCast(c_int, machine_idx), &p, &n));
    if (n == 0) return &[_]VAD.VADSpeechSegment{};
    return @ptrCast([*]const VAD.VADSpeechSegment, p)[0..n];
}

pub fn durationToSamples(sample_rate: usize, buffer_duration: f32) usize { // src/AudioPipeline.zig:146-149
    const duration_f = (@intToFloat(f32, sample_rate) * buffer_duration);
    return @floatToInt(usize, @ceil(duration_f));
}
