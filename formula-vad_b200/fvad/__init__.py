"""Python host side of libfvad_cuda.so (ctypes over the C ABI in include/fvad.h).

Mirrors the reference's operator interface for the hot path -- AudioPipeline
(src/AudioPipeline.zig), FFT (src/FFT.zig), Denoiser (src/Denoiser.zig) -- with the same
names, argument meaning and error names.  There is no CPU fallback: importing works
anywhere, but every compute call raises FvadError(CudaError) without a B200.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libfvad_cuda.so")

TRACE_RECORDS, TRACE_DENOISED, TRACE_RNN_VAD, TRACE_SPECTRA = 1, 2, 4, 8
KEEP_FEATURES = 16        # FVAD_KEEP_FEATURES: per-STFT-frame features stay in HBM for PipelineBatch.sweep


class VadMachineConfig(C.Structure):
    """VADMachine.Config (src/AudioPipeline/VADMachine.zig:18-39)."""
    _fields_ = [("speech_min_freq", C.c_float), ("speech_max_freq", C.c_float),
                ("long_term_speech_avg_sec", C.c_float), ("has_initial_long_term_avg", C.c_int32),
                ("initial_long_term_avg", C.c_double), ("short_term_speech_avg_sec", C.c_float),
                ("speech_threshold_factor", C.c_float), ("channel_vol_ratio_avg_sec", C.c_float),
                ("channel_vol_ratio_threshold", C.c_float), ("min_consecutive_sec_to_open", C.c_float),
                ("max_speech_gap_sec", C.c_float), ("min_vad_duration_sec", C.c_float)]


class VadConfig(C.Structure):
    """VAD.Config (src/AudioPipeline/VAD.zig:17-23)."""
    _fields_ = [("fft_size", C.c_size_t), ("use_denoiser", C.c_int32), ("vad_machine_config", VadMachineConfig),
                ("alt_vad_machine_configs", C.POINTER(VadMachineConfig)), ("n_alt_vad_machine_configs", C.c_size_t)]


class Config(C.Structure):
    """AudioPipeline.Config (src/AudioPipeline.zig:20-26)."""
    _fields_ = [("sample_rate", C.c_size_t), ("n_channels", C.c_size_t), ("buffer_length", C.c_size_t),
                ("vad_config", VadConfig), ("skip_processing", C.c_int32)]


SEG_DTYPE = np.dtype([("sample_from", "<u8"), ("sample_to", "<u8"), ("rnn_vad", "<f4"), ("vol_ratio", "<f4")])
REC_DTYPE = np.dtype([("index", "<u8"), ("min_volume", "<f4"), ("vol_ratio", "<f4"), ("rnn_vad", "<f4"), ("has_vad", "<f4")])
EVENT_DTYPE = np.dtype([("at_index", "<u8"), ("recording_state", "<i4"), ("_pad", "<i4"), ("sample_number", "<u8")])

_fp = C.POINTER(C.c_float)
_lib = None


class Clip(C.Structure):
    """fvad_clip = the AudioBuffer Callbacks.on_recording receives (src/AudioPipeline.zig:15-18, Recorder.zig:121-146)."""
    _fields_ = [("global_start_frame_number", C.c_uint64), ("length", C.c_size_t), ("sample_rate", C.c_size_t),
                ("channel", C.c_size_t), ("duration_seconds", C.c_float), ("lost", C.c_int32), ("pcm", C.POINTER(C.c_float))]


ON_RECORDING = C.CFUNCTYPE(None, C.c_void_p, C.c_size_t, C.POINTER(Clip))


def _clip_dict(c):
    pcm = None if c.lost or not c.pcm else np.ctypeslib.as_array(c.pcm, shape=(c.length,)).copy()
    return dict(start=int(c.global_start_frame_number), end=int(c.global_start_frame_number + c.length), length=int(c.length),
                sample_rate=int(c.sample_rate), channel=int(c.channel), duration_seconds=float(c.duration_seconds),
                lost=bool(c.lost), pcm=pcm)


class SpectrogramInfo(C.Structure):
    _fields_ = [("height", C.c_size_t), ("width", C.c_size_t), ("bin_width", C.c_float), ("nyquist_freq", C.c_float),
                ("length_sec", C.c_double)]


class FvadError(RuntimeError):
    """Carries the reference's Zig error name (e.g. 'InvalidSampleRate') as .name."""

    def __init__(self, status, message):
        self.status = status
        self.name = lib().fvad_status_name(status).decode()
        super().__init__(f"{self.name} ({status}): {message}")


def lib():
    """Load libfvad_cuda.so; fails loudly if the CUDA extension was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, sz, u64p = C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)
    szp = C.POINTER(C.c_size_t)
    sig = {
        "fvad_last_error": (C.c_char_p, []), "fvad_status_name": (C.c_char_p, [C.c_int]),
        "fvad_version": (C.c_char_p, []), "fvad_device_count": (C.c_int, []), "fvad_launch_count": (C.c_uint64, []),
        "fvad_debug_set": (C.c_int, [C.c_char_p, C.c_char_p]),
        "fvad_debug_log10": (C.c_int, [_fp, _fp, sz]),
        "fvad_profile_enable": (None, [C.c_int]), "fvad_profile_report": (C.c_int, [C.c_char_p, sz]),
        "fvad_default_vad_machine_config": (None, [C.POINTER(VadMachineConfig)]),
        "fvad_default_config": (None, [C.POINTER(Config), sz, sz]),
        "fvad_batch_create": (C.c_int, [C.POINTER(Config), sz, C.c_int, C.POINTER(vp)]),
        "fvad_batch_destroy": (None, [vp]),
        "fvad_batch_push": (C.c_int, [vp, C.POINTER(_fp), sz, u64p]),
        "fvad_batch_push_device": (C.c_int, [vp, vp, sz, sz, sz, u64p]),
        "fvad_batch_sync": (C.c_int, [vp]), "fvad_batch_reset": (C.c_int, [vp]),
        "fvad_batch_segments": (C.c_int, [vp, sz, C.c_int, C.POINTER(vp), szp]),
        "fvad_batch_events": (C.c_int, [vp, sz, C.POINTER(vp), szp]),
        "fvad_batch_counts": (C.c_int, [vp, u64p, u64p]),
        "fvad_batch_cuda_stream": (vp, [vp]),
        "fvad_batch_set_recording": (C.c_int, [vp, sz, ON_RECORDING, vp]),
        "fvad_batch_clips": (C.c_int, [vp, sz, C.POINTER(C.POINTER(Clip)), szp]),
        "fvad_batch_sweep": (C.c_int, [vp, C.POINTER(VadMachineConfig), sz, sz, vp, vp, C.c_int]),
        "fvad_batch_feature_count": (C.c_int, [vp, szp]),
        "fvad_batch_features": (C.c_int, [vp, sz, vp, sz, szp]),
        "fvad_batch_set_trace": (C.c_int, [vp, C.c_uint]),
        "fvad_batch_trace_records": (C.c_int, [vp, sz, vp, sz, szp]),
        "fvad_batch_trace_denoised": (C.c_int, [vp, sz, sz, _fp, sz, szp]),
        "fvad_batch_trace_rnn_vad": (C.c_int, [vp, sz, _fp, sz, szp]),
        "fvad_batch_trace_spectra": (C.c_int, [vp, sz, sz, _fp, sz, szp]),
        "fvad_pipeline_create": (C.c_int, [C.POINTER(Config), C.POINTER(vp)]),
        "fvad_pipeline_push": (C.c_int, [vp, C.POINTER(_fp), sz, u64p]),
        "fvad_pipeline_segments": (C.c_int, [vp, C.c_int, C.POINTER(vp), szp]),
        "fvad_pipeline_destroy": (None, [vp]),
        "fvad_fft_create": (C.c_int, [sz, sz, C.POINTER(vp)]), "fvad_fft_destroy": (None, [vp]),
        "fvad_fft_run": (C.c_int, [vp, _fp, sz, _fp, sz, _fp, sz, _fp, sz]),
        "fvad_fft_bin_count": (sz, [vp]), "fvad_fft_bin_width": (C.c_float, [vp]),
        "fvad_fft_nyquist_freq": (C.c_float, [vp]),
        "fvad_fft_freq_to_bin": (C.c_int, [vp, C.c_float, szp]),
        "fvad_fft_bin_to_freq": (C.c_int, [vp, sz, _fp]),
        "fvad_hann_window_periodic": (None, [_fp, sz]), "fvad_window_norm_factor": (C.c_float, [_fp, sz]),
        "fvad_stft_batch": (C.c_int, [vp, vp, sz, sz, sz, sz, sz, sz, vp, vp, vp]),
        "fvad_spectrogram_shape": (C.c_int, [sz, sz, sz, sz, C.POINTER(SpectrogramInfo)]),
        "fvad_spectrogram_compute": (C.c_int, [_fp, sz, sz, sz, sz, _fp, sz, _fp, C.POINTER(SpectrogramInfo)]),
        "fvad_denoiser_create": (C.c_int, [C.POINTER(vp)]), "fvad_denoiser_destroy": (None, [vp]),
        "fvad_denoiser_frame_size": (sz, []),
        "fvad_denoiser_denoise": (C.c_int, [vp, _fp, sz, _fp, sz, _fp, sz, _fp]),
        "fvad_denoiser_batch_create": (C.c_int, [sz, C.c_int, C.POINTER(vp)]),
        "fvad_denoiser_batch_destroy": (None, [vp]),
        "fvad_denoiser_batch_run": (C.c_int, [vp, C.POINTER(_fp), sz, C.POINTER(_fp), C.POINTER(_fp)]),
        "fvad_vadm_sweep": (C.c_int, [vp, sz, sz, C.POINTER(VadMachineConfig), sz, sz, sz, C.c_int, vp, sz, vp]),
        "fvad_vadm_sweep_static": (C.c_int, [vp, vp, vp, sz, sz, C.c_uint64, C.POINTER(VadMachineConfig), sz, sz, sz, vp, sz, vp, vp]),
        "fvad_eval_stats_batch": (C.c_int, [vp, sz, sz, vp, vp, vp, vp, vp, sz, sz, vp, C.c_int, vp]),
        "fvad_eval_stats_batch_device": (C.c_int, [vp, sz, sz, vp, vp, vp, vp, vp, sz, sz, vp, vp]),
        "fvad_synth_fill_host": (C.c_int, [_fp, C.c_uint64, C.c_uint64, C.c_uint64, sz]),
        "fvad_synth_fill_device": (C.c_int, [vp, sz, C.c_uint64, C.c_uint64, sz, C.c_uint64, sz, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here = header/library mismatch: fail loudly
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise FvadError(rc, lib().fvad_last_error().decode())


def _f(a):
    return a.ctypes.data_as(_fp)


def default_machine_config(**kw):
    c = VadMachineConfig()
    lib().fvad_default_vad_machine_config(C.byref(c))
    for k, v in kw.items():
        if k == "initial_long_term_avg":
            if v is None:
                c.has_initial_long_term_avg = 0
            else:
                c.has_initial_long_term_avg, c.initial_long_term_avg = 1, v
        else:
            setattr(c, k, v)
    return c


def debug_set(key=None, value=None):
    """fvad_debug_set: measurement / test knobs (kernel shapes, tile length, ...).  value None clears the key, key None
    clears all.  The library reads no environment variables."""
    _check(lib().fvad_debug_set(None if key is None else str(key).encode(), None if value is None else str(value).encode()))


def debug_log10(ex):
    """The device's (float)log10(1e-2 + ex) of rnnoise's feature extraction, elementwise (fvad_debug_log10)."""
    ex = np.ascontiguousarray(ex, np.float32)
    out = np.empty_like(ex)
    _check(lib().fvad_debug_log10(_f(ex), _f(out), ex.size))
    return out


def launch_count():
    return int(lib().fvad_launch_count())


def profile_enable(on=True):
    lib().fvad_profile_enable(int(on))


def profile_report():
    import json
    n = lib().fvad_profile_report(None, 0)
    buf = C.create_string_buffer(n + 16)
    lib().fvad_profile_report(buf, n + 16)   # the size query above leaves the records in place
    return json.loads(buf.value.decode() or "{}")


class PipelineBatch:
    """n_streams independent AudioPipelines sharing one Config, on one GPU."""

    def __init__(self, n_streams, sample_rate=48000, n_channels=1, fft_size=2048, use_denoiser=True,
                 vad_machine_config=None, alt_vad_machine_configs=(), skip_processing=False, device=-1, trace=0):
        L = lib()
        cfg = Config()
        L.fvad_default_config(C.byref(cfg), sample_rate, n_channels)
        cfg.vad_config.fft_size = fft_size
        cfg.vad_config.use_denoiser = int(use_denoiser)
        if vad_machine_config is not None:
            cfg.vad_config.vad_machine_config = vad_machine_config
        self._alts = (VadMachineConfig * max(1, len(alt_vad_machine_configs)))(*alt_vad_machine_configs)
        cfg.vad_config.alt_vad_machine_configs = self._alts
        cfg.vad_config.n_alt_vad_machine_configs = len(alt_vad_machine_configs)
        cfg.skip_processing = int(skip_processing)
        self.n_streams, self.n_channels, self.n_machines = n_streams, n_channels, 1 + len(alt_vad_machine_configs)
        self.fft_size = fft_size
        self._h = C.c_void_p()
        _check(L.fvad_batch_create(C.byref(cfg), n_streams, device, C.byref(self._h)))
        if trace:
            _check(L.fvad_batch_set_trace(self._h, trace))

    def close(self):
        if getattr(self, "_h", None) and self._h:
            lib().fvad_batch_destroy(self._h)
            self._h = None

    deinit = close

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def push_samples(self, pcm):
        """pcm: array [n_streams][n_channels][n] (or [n_streams][n] for mono) of host f32.  Returns the
        first sample index (AudioPipeline.pushSamples, src/AudioPipeline.zig:86-120)."""
        a = np.asarray(pcm, dtype=np.float32)
        if a.ndim == 2:
            a = a[:, None, :]
        if a.shape[0] != self.n_streams or a.shape[1] != self.n_channels:
            raise FvadError(-11, f"expected [{self.n_streams}][{self.n_channels}][n], got {a.shape}")
        a = np.ascontiguousarray(a)
        n = a.shape[2]
        ptrs = (_fp * (self.n_streams * self.n_channels))()
        base = a.ctypes.data
        for r in range(self.n_streams * self.n_channels):
            ptrs[r] = C.cast(base + 4 * n * r, _fp)
        first = C.c_uint64()
        _check(lib().fvad_batch_push(self._h, ptrs, n, C.byref(first)))
        return first.value

    def push_rows(self, ptrs, n):
        """Raw form: ptrs = list of host addresses (ints), one per (stream, channel) row."""
        arr = (_fp * len(ptrs))(*[C.cast(p, _fp) for p in ptrs])
        first = C.c_uint64()
        _check(lib().fvad_batch_push(self._h, arr, n, C.byref(first)))
        return first.value

    def push_device(self, dev_ptr, stream_stride, channel_stride, n):
        first = C.c_uint64()
        _check(lib().fvad_batch_push_device(self._h, C.c_void_p(dev_ptr), stream_stride, channel_stride, n, C.byref(first)))
        return first.value

    def sync(self):
        _check(lib().fvad_batch_sync(self._h))

    def cuda_stream(self):
        """Address of the batch's main cudaStream_t (for torch.cuda.ExternalStream / event timing)."""
        return int(lib().fvad_batch_cuda_stream(self._h) or 0)

    def reset(self):
        _check(lib().fvad_batch_reset(self._h))

    def set_recording(self, history_samples=0, on_recording=None):
        """Recorder / Callbacks.on_recording through the C ABI (fvad_batch_set_recording): before the first push.
        history_samples = 0 keeps every pushed sample in HBM (lossless); else a ring of that many samples per channel.
        on_recording(stream, clip_dict) is called from inside push / sync, once per clip."""
        if on_recording is None:
            self._rec_cb = C.cast(None, ON_RECORDING)
        else:
            def tramp(_ctx, stream, clip):
                on_recording(int(stream), _clip_dict(clip.contents))
            self._rec_cb = ON_RECORDING(tramp)                  # kept alive with the batch
        _check(lib().fvad_batch_set_recording(self._h, history_samples, self._rec_cb, None))

    def clips(self, stream):
        """fvad_batch_clips: the clips completed since the previous call for this stream, as dicts (pcm copied)."""
        p, n = C.POINTER(Clip)(), C.c_size_t()
        _check(lib().fvad_batch_clips(self._h, stream, C.byref(p), C.byref(n)))
        return [_clip_dict(p[i]) for i in range(n.value)]

    def feature_count(self):
        n = C.c_size_t()
        _check(lib().fvad_batch_feature_count(self._h, C.byref(n)))
        return n.value

    def features(self, stream):
        """The kept features of one stream as REC_DTYPE records (trace=KEEP_FEATURES)."""
        n = C.c_size_t()
        _check(lib().fvad_batch_features(self._h, stream, None, 0, C.byref(n)))
        out = np.zeros(n.value, REC_DTYPE)
        _check(lib().fvad_batch_features(self._h, stream, out.ctypes.data, n.value, C.byref(n)))
        return out

    def sweep(self, configs, max_segs=512, d_segs=0, d_counts=0):
        """fvad_batch_sweep: further VADMachine configs over the features the batch kept (trace=KEEP_FEATURES), on the
        device.  Host form (default): returns [stream][config] -> SEG_DTYPE arrays (re-run with a larger max_segs if a
        list was truncated).  Device form: d_segs / d_counts are device addresses of [S*K*max_segs] fvad_segment and
        [S*K] u32; returns None, the lists stay on the device (fvad.evaluator.stats_batch_device consumes them)."""
        K = len(configs)
        cfgs = (VadMachineConfig * K)(*configs)
        if d_segs:
            _check(lib().fvad_batch_sweep(self._h, cfgs, K, max_segs, C.c_void_p(d_segs), C.c_void_p(d_counts), 1))
            return None
        segs = np.zeros((self.n_streams * K, max_segs), SEG_DTYPE)
        counts = np.zeros(self.n_streams * K, np.uint32)
        _check(lib().fvad_batch_sweep(self._h, cfgs, K, max_segs, segs.ctypes.data, counts.ctypes.data, 0))
        if counts.max(initial=0) > max_segs:
            return self.sweep(configs, int(counts.max()))
        return [[segs[s * K + k, :counts[s * K + k]].copy() for k in range(K)] for s in range(self.n_streams)]

    def counts(self):
        a, b = C.c_uint64(), C.c_uint64()
        _check(lib().fvad_batch_counts(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def segments(self, stream, machine_idx=0):
        p, n = C.c_void_p(), C.c_size_t()
        _check(lib().fvad_batch_segments(self._h, stream, machine_idx, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros(0, SEG_DTYPE)
        buf = (C.c_char * (n.value * SEG_DTYPE.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=SEG_DTYPE).copy()

    def events(self, stream):
        p, n = C.c_void_p(), C.c_size_t()
        _check(lib().fvad_batch_events(self._h, stream, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros(0, EVENT_DTYPE)
        buf = (C.c_char * (n.value * EVENT_DTYPE.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=EVENT_DTYPE).copy()

    def trace_records(self, stream):
        n = C.c_size_t()
        _check(lib().fvad_batch_trace_records(self._h, stream, None, 0, C.byref(n)))
        out = np.zeros(n.value, REC_DTYPE)
        _check(lib().fvad_batch_trace_records(self._h, stream, out.ctypes.data, n.value, C.byref(n)))
        return out

    def _trace_f(self, fn, *idx):
        n = C.c_size_t()
        _check(fn(self._h, *idx, None, 0, C.byref(n)))
        out = np.zeros(n.value, np.float32)
        _check(fn(self._h, *idx, _f(out), n.value, C.byref(n)))
        return out

    def trace_denoised(self, stream, channel=0):
        return self._trace_f(lib().fvad_batch_trace_denoised, stream, channel)

    def trace_rnn_vad(self, stream):
        return self._trace_f(lib().fvad_batch_trace_rnn_vad, stream)

    def trace_spectra(self, stream, channel=0):
        return self._trace_f(lib().fvad_batch_trace_spectra, stream, channel).reshape(-1, self.fft_size // 2 + 1)


class AudioPipeline(PipelineBatch):
    """One pipeline: AudioPipeline.init / pushSamples / deinit, vad_segments read-out."""

    def __init__(self, sample_rate=48000, n_channels=1, **kw):
        super().__init__(1, sample_rate=sample_rate, n_channels=n_channels, **kw)

    def push_samples(self, channel_pcm):
        a = np.asarray(channel_pcm, dtype=np.float32)
        if a.ndim == 1:
            a = a[None, :]
        return super().push_samples(a[None, :, :])

    @property
    def vad_segments(self):
        return self.segments(0, 0)


class FFT:
    """FFT.zig drop-in: init(n_fft, sample_rate), fft(samples, window) -> normalised magnitudes."""

    def __init__(self, n_fft, sample_rate):
        self._h = C.c_void_p()
        _check(lib().fvad_fft_create(n_fft, sample_rate, C.byref(self._h)))
        self.n_fft, self.sample_rate = n_fft, sample_rate

    def deinit(self):
        if self._h:
            lib().fvad_fft_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.deinit()
        except Exception:
            pass

    def fft(self, samples, window, second=None, result_len=None):
        first = np.ascontiguousarray(samples, np.float32)
        second = np.zeros(0, np.float32) if second is None else np.ascontiguousarray(second, np.float32)
        window = np.ascontiguousarray(window, np.float32)
        res = np.empty(self.bin_count() if result_len is None else result_len, np.float32)
        _check(lib().fvad_fft_run(self._h, _f(first), first.size, _f(second), second.size, _f(window), window.size,
                                  _f(res), res.size))
        return res

    def bin_count(self):
        return int(lib().fvad_fft_bin_count(self._h))

    def bin_width(self):
        return float(lib().fvad_fft_bin_width(self._h))

    def nyquist_freq(self):
        return float(lib().fvad_fft_nyquist_freq(self._h))

    def freq_to_bin(self, freq):
        b = C.c_size_t()
        _check(lib().fvad_fft_freq_to_bin(self._h, freq, C.byref(b)))
        return b.value

    def bin_to_freq(self, b):
        f = C.c_float()
        _check(lib().fvad_fft_bin_to_freq(self._h, b, C.byref(f)))
        return f.value

    def stft_batch(self, dev_ptr, n_rows, row_stride, n_frames, hop, min_bin, max_bin, band_ptr, spectra_ptr=0,
                   cuda_stream=0):
        _check(lib().fvad_stft_batch(self._h, C.c_void_p(dev_ptr), n_rows, row_stride, n_frames, hop, min_bin, max_bin,
                                     C.c_void_p(band_ptr), C.c_void_p(spectra_ptr) if spectra_ptr else None,
                                     C.c_void_p(cuda_stream) if cuda_stream else None))


class Spectrogram:
    """audio_utils/Spectrogram.zig drop-in: Spectrogram.compute(samples, sample_rate, fft_size, hop_size) with the
    reference's fields values (width x height, row = one FFT step), bin_labels, bin_width, height, width,
    nyquist_freq, length_sec.  Errors carry the Zig names (InsufficientSamples, ...)."""

    @staticmethod
    def shape(n_samples, sample_rate, fft_size=2048, hop_size=2048):
        info = SpectrogramInfo()
        _check(lib().fvad_spectrogram_shape(n_samples, sample_rate, fft_size, hop_size, C.byref(info)))
        return info

    @classmethod
    def compute(cls, samples, sample_rate, fft_size=2048, hop_size=2048):
        samples = np.ascontiguousarray(samples, np.float32)
        info = cls.shape(samples.size, sample_rate, fft_size, hop_size)
        self = cls()
        self.values = np.empty((info.width, info.height), np.float32)
        self.bin_labels = np.empty(info.height, np.float32)
        _check(lib().fvad_spectrogram_compute(_f(samples), samples.size, sample_rate, fft_size, hop_size, _f(self.values),
                                              self.values.size, _f(self.bin_labels), C.byref(info)))
        self.height, self.width = int(info.height), int(info.width)
        self.bin_width, self.nyquist_freq, self.length_sec = float(info.bin_width), float(info.nyquist_freq), float(info.length_sec)
        return self


def norm_to_dbfs(values):
    """audio_utils.normToDBFS (src/audio_utils.zig:6-12): 20*log10(v), host-side (numpy's f32 log10 may differ from
    Zig's by an ulp; it feeds only the viewer)."""
    return (np.float32(20.0) * np.log10(np.asarray(values, np.float32))).astype(np.float32)


def hann_window_periodic(n):
    w = np.empty(n, np.float32)
    lib().fvad_hann_window_periodic(_f(w), n)
    return w


class Denoiser:
    """Denoiser.zig drop-in: init(), denoise(samples) -> (denoised, vad), getFrameSize()."""

    def __init__(self):
        self._h = C.c_void_p()
        _check(lib().fvad_denoiser_create(C.byref(self._h)))

    def deinit(self):
        if self._h:
            lib().fvad_denoiser_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.deinit()
        except Exception:
            pass

    @staticmethod
    def get_frame_size():
        return int(lib().fvad_denoiser_frame_size())

    def denoise(self, samples, second=None, result_len=480):
        first = np.ascontiguousarray(samples, np.float32)
        second = np.zeros(0, np.float32) if second is None else np.ascontiguousarray(second, np.float32)
        res = np.empty(result_len, np.float32)
        vad = C.c_float()
        _check(lib().fvad_denoiser_denoise(self._h, _f(first), first.size, _f(second), second.size, _f(res), res.size,
                                           C.byref(vad)))
        return res, vad.value


class DenoiserBatch:
    """Many independent Denoisers x many consecutive frames per call."""

    def __init__(self, n_streams, device=-1):
        self._h = C.c_void_p()
        self.n_streams = n_streams
        _check(lib().fvad_denoiser_batch_create(n_streams, device, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().fvad_denoiser_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run(self, pcm):
        a = np.ascontiguousarray(pcm, np.float32)
        assert a.ndim == 2 and a.shape[0] == self.n_streams and a.shape[1] % 480 == 0
        nf = a.shape[1] // 480
        out = np.empty_like(a)
        vad = np.empty((self.n_streams, nf), np.float32)
        pi, po, pv = (_fp * self.n_streams)(), (_fp * self.n_streams)(), (_fp * self.n_streams)()
        for s in range(self.n_streams):
            pi[s] = C.cast(a.ctypes.data + 4 * a.shape[1] * s, _fp)
            po[s] = C.cast(out.ctypes.data + 4 * a.shape[1] * s, _fp)
            pv[s] = C.cast(vad.ctypes.data + 4 * nf * s, _fp)
        _check(lib().fvad_denoiser_batch_run(self._h, pi, nf, po, pv))
        return out, vad


def vadm_sweep(records, configs, sample_rate=48000, fft_size=2048, device=-1, max_segs=512):
    """records: REC_DTYPE array [n_streams][n_records]; configs: list of VadMachineConfig.
    Returns a list [stream][config] of SEG_DTYPE arrays."""
    rec = np.ascontiguousarray(records, REC_DTYPE)
    if rec.ndim == 1:
        rec = rec[None, :]
    S, N = rec.shape
    K = len(configs)
    cfgs = (VadMachineConfig * K)(*configs)
    segs = np.zeros((S * K, max_segs), SEG_DTYPE)
    counts = np.zeros(S * K, np.uint32)
    _check(lib().fvad_vadm_sweep(rec.ctypes.data, S, N, cfgs, K, sample_rate, fft_size, device, segs.ctypes.data, max_segs,
                                 counts.ctypes.data))
    if counts.max(initial=0) > max_segs:
        return vadm_sweep(records, configs, sample_rate, fft_size, device, int(counts.max()))
    return [[segs[s * K + k, :counts[s * K + k]].copy() for k in range(K)] for s in range(S)]


VADM_SWEEP_STATIC_SCRATCH_BYTES = 16 << 20      # FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES
VADM_SWEEP_STATIC_MAX_CONFIGS = 256             # FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS


def vadm_sweep_static(d_min_volume, d_vol_ratio, d_rnn_vad, n_streams, n_records, first_index, configs, d_segs, max_segs,
                      d_counts, sample_rate=48000, fft_size=2048, cuda_stream=0):
    """fvad_vadm_sweep_static: the sweep on DEVICE planes into DEVICE outputs with no allocation (scratch is a static
    device array of VADM_SWEEP_STATIC_SCRATCH_BYTES).  All pointer arguments are device addresses (ints); the call is
    asynchronous on `cuda_stream` of the current device.  No reference counterpart."""
    K = len(configs)
    cfgs = (VadMachineConfig * K)(*configs)
    _check(lib().fvad_vadm_sweep_static(C.c_void_p(d_min_volume), C.c_void_p(d_vol_ratio),
                                        C.c_void_p(d_rnn_vad) if d_rnn_vad else None, n_streams, n_records, first_index,
                                        cfgs, K, sample_rate, fft_size, C.c_void_p(d_segs), max_segs, C.c_void_p(d_counts),
                                        C.c_void_p(cuda_stream) if cuda_stream else None))


def synth_host(seed, stream_id, first_sample, n):
    out = np.empty(n, np.float32)
    _check(lib().fvad_synth_fill_host(_f(out), seed, stream_id, first_sample, n))
    return out


def synth_device(dev_ptr, stream_stride, seed, first_stream_id, n_streams, first_sample, n, cuda_stream=0):
    _check(lib().fvad_synth_fill_device(C.c_void_p(dev_ptr), stream_stride, seed, first_stream_id, n_streams, first_sample,
                                        n, C.c_void_p(cuda_stream) if cuda_stream else None))
