"""ctypes front-end of the parity oracle.  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product (formula-vad_b200/).  See oracle/oracle.h."""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_port.so")
REF_SO = os.path.join(HERE, "_ref", "liboracle_ref.so")
REF_FAST_SO = os.path.join(HERE, "_ref", "liboracle_ref_fast.so")      # timing stand-in for zig ReleaseFast (Makefile: ref_fast)
SYNTH_SO = os.path.join(HERE, "libsynth_host.so")


class VadmConfig(C.Structure):
    _fields_ = [("speech_min_freq", C.c_float), ("speech_max_freq", C.c_float),
                ("long_term_speech_avg_sec", C.c_float), ("has_initial_long_term_avg", C.c_int),
                ("initial_long_term_avg", C.c_double), ("short_term_speech_avg_sec", C.c_float),
                ("speech_threshold_factor", C.c_float), ("channel_vol_ratio_avg_sec", C.c_float),
                ("channel_vol_ratio_threshold", C.c_float), ("min_consecutive_sec_to_open", C.c_float),
                ("max_speech_gap_sec", C.c_float), ("min_vad_duration_sec", C.c_float)]


class Segment(C.Structure):
    _fields_ = [("sample_from", C.c_uint64), ("sample_to", C.c_uint64),
                ("debug_rnn_vad", C.c_float), ("debug_avg_speech_vol_ratio", C.c_float)]


class StftRecord(C.Structure):
    _fields_ = [("index", C.c_uint64), ("min_volume", C.c_float), ("vol_ratio", C.c_float),
                ("rnn_vad", C.c_float), ("has_vad", C.c_float)]


class CaptureEvent(C.Structure):
    _fields_ = [("at_index", C.c_uint64), ("kind", C.c_int), ("sample_number", C.c_uint64)]


class RnnoiseTrace(C.Structure):
    _fields_ = [("hp", C.c_float * 480), ("X", C.c_float * 962), ("Ex", C.c_float * 22), ("Ep", C.c_float * 22),
                ("Exp", C.c_float * 22), ("pitch_index", C.c_int), ("pitch_gain", C.c_float), ("silence", C.c_int),
                ("features", C.c_float * 42), ("gains", C.c_float * 22), ("vad_prob", C.c_float),
                ("out", C.c_float * 480)]


SEG_DTYPE = np.dtype([("sample_from", "<u8"), ("sample_to", "<u8"), ("rnn_vad", "<f4"), ("vol_ratio", "<f4")])
REC_DTYPE = np.dtype([("index", "<u8"), ("min_volume", "<f4"), ("vol_ratio", "<f4"), ("rnn_vad", "<f4"), ("has_vad", "<f4")])

_fp = C.POINTER(C.c_float)


def _f(a):
    return a.ctypes.data_as(_fp)


def build(kind="port"):
    target = "port" if kind == "port" else "ref"
    subprocess.run(["make", "-s", "-C", HERE, target], check=True)


class Oracle:
    """One loaded oracle build: 'port', 'reference', or 'reference_fast' (the reference's sources at -O3 with AVX2/FMA
    and contraction allowed: for TIMING only, never a parity checker)."""

    def __init__(self, kind="port"):
        path = {"port": PORT_SO, "reference": REF_SO, "reference_fast": REF_FAST_SO}[kind]
        if not os.path.exists(path):
            build("port" if kind == "port" else "reference")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.kind = kind
        kind = "reference" if kind == "reference_fast" else kind
        L = self.lib = C.CDLL(path)
        L.orc_build_kind.restype = C.c_char_p
        L.orc_fft_create.restype = C.c_void_p
        L.orc_fft_create.argtypes = [C.c_int, C.c_int, C.c_int]
        L.orc_fft_destroy.argtypes = [C.c_void_p]
        L.orc_fft_run.argtypes = [C.c_void_p, _fp, C.c_size_t, _fp, C.c_size_t, _fp, C.c_size_t, _fp, C.c_size_t]
        L.orc_fft_last_spectrum.restype = _fp
        L.orc_fft_last_spectrum.argtypes = [C.c_void_p]
        L.orc_fft_freq_to_bin.restype = C.c_long
        L.orc_fft_freq_to_bin.argtypes = [C.c_void_p, C.c_float]
        L.orc_hann_window_periodic.argtypes = [_fp, C.c_size_t]
        L.orc_window_norm_factor.restype = C.c_float
        L.orc_window_norm_factor.argtypes = [_fp, C.c_size_t]
        L.orc_rms_volume.restype = C.c_float
        L.orc_rms_volume.argtypes = [_fp, C.c_size_t, _fp, C.c_size_t]
        L.orc_denoiser_create.restype = C.c_void_p
        L.orc_denoiser_destroy.argtypes = [C.c_void_p]
        L.orc_denoiser_denoise.restype = C.c_float
        L.orc_denoiser_denoise.argtypes = [C.c_void_p, _fp, _fp]
        L.orc_denoiser_process_raw.restype = C.c_float
        L.orc_denoiser_process_raw.argtypes = [C.c_void_p, _fp, _fp, C.POINTER(RnnoiseTrace)]
        L.orc_pipeline_create.restype = C.c_void_p
        L.orc_pipeline_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(VadmConfig),
                                          C.POINTER(VadmConfig), C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.orc_pipeline_destroy.argtypes = [C.c_void_p]
        L.orc_pipeline_push.restype = C.c_uint64
        L.orc_pipeline_push.argtypes = [C.c_void_p, C.POINTER(_fp), C.c_size_t]
        L.orc_pipeline_segments.restype = C.c_size_t
        L.orc_pipeline_segments.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.orc_pipeline_events.restype = C.c_size_t
        L.orc_pipeline_events.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_pipeline_enable_trace.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_pipeline_enable_spectra.argtypes = [C.c_void_p, C.c_int]
        L.orc_pipeline_stft_records.restype = C.c_size_t
        L.orc_pipeline_stft_records.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_pipeline_denoised.restype = C.c_size_t
        L.orc_pipeline_denoised.argtypes = [C.c_void_p, C.c_int, _fp, C.c_size_t]
        L.orc_pipeline_rnn_vad.restype = C.c_size_t
        L.orc_pipeline_rnn_vad.argtypes = [C.c_void_p, _fp, C.c_size_t]
        L.orc_pipeline_spectra.restype = C.c_size_t
        L.orc_pipeline_spectra.argtypes = [C.c_void_p, C.c_int, _fp, C.c_size_t]
        L.orc_vadm_scan.restype = C.c_size_t
        L.orc_vadm_scan.argtypes = [C.POINTER(VadmConfig), C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.orc_default_vadm_config.argtypes = [C.POINTER(VadmConfig)]
        assert L.orc_build_kind().decode() == kind, (L.orc_build_kind(), kind)

    # -- helpers -----------------------------------------------------------
    def default_config(self, **kw):
        c = VadmConfig()
        self.lib.orc_default_vadm_config(C.byref(c))
        for k, v in kw.items():
            if k == "initial_long_term_avg":
                if v is None:
                    c.has_initial_long_term_avg = 0
                else:
                    c.has_initial_long_term_avg, c.initial_long_term_avg = 1, v
            else:
                setattr(c, k, v)
        return c

    def hann(self, n):
        w = np.empty(n, np.float32)
        self.lib.orc_hann_window_periodic(_f(w), n)
        return w

    def fft(self, samples, window=None, n_fft=None, faithful=True, split=None, want_spectrum=False):
        """FFT.fft: returns normalised magnitudes (n_fft/2+1)."""
        samples = np.ascontiguousarray(samples, np.float32)
        n_fft = n_fft or samples.size
        if window is None:
            window = self.hann(n_fft)
        h = self.lib.orc_fft_create(n_fft, 48000, int(faithful))
        if not h:
            raise ValueError("InvalidFFTSize")
        try:
            res = np.empty(n_fft // 2 + 1, np.float32)
            k = samples.size if split is None else split
            a, b = samples[:k].copy(), samples[k:].copy()
            rc = self.lib.orc_fft_run(h, _f(a), a.size, _f(b), b.size, _f(window), window.size, _f(res), res.size)
            if rc:
                raise ValueError(f"orc_fft_run -> {rc}")
            if want_spectrum:
                sp = np.ctypeslib.as_array(self.lib.orc_fft_last_spectrum(h), shape=(n_fft // 2 + 1, 2)).copy()
                return res, sp
            return res
        finally:
            self.lib.orc_fft_destroy(h)

    def denoise_stream(self, pcm, raw=False, trace=False):
        """Run Denoiser.denoise over consecutive 480-sample frames of a mono signal.
        Returns (denoised, vad[, traces])."""
        pcm = np.ascontiguousarray(pcm, np.float32)
        nf = pcm.size // 480
        out = np.empty(nf * 480, np.float32)
        vad = np.empty(nf, np.float32)
        d = self.lib.orc_denoiser_create()
        traces = []
        try:
            for q in range(nf):
                i = pcm[q * 480:(q + 1) * 480]
                o = out[q * 480:(q + 1) * 480]
                if raw or trace:
                    t = RnnoiseTrace() if trace else None
                    vad[q] = self.lib.orc_denoiser_process_raw(d, _f(i), _f(o), C.byref(t) if trace else None)
                    if trace:
                        traces.append(t)
                else:
                    vad[q] = self.lib.orc_denoiser_denoise(d, _f(i), _f(o))
        finally:
            self.lib.orc_denoiser_destroy(d)
        return (out, vad, traces) if trace else (out, vad)

    def run_pipeline(self, channels, use_denoiser=True, fft_size=2048, main_cfg=None, alt_cfgs=(), faithful=True,
                     push=48000, trace=False, spectra=False, sample_rate=48000):
        """AudioPipeline over planar channels [C][n]; pushes `push` samples at a time like
        SimulationInstance.simulateVAD (SimulationInstance.zig:194-203)."""
        ch = [np.ascontiguousarray(c, np.float32) for c in channels]
        n = ch[0].size
        main_cfg = main_cfg or self.default_config()
        alts = (VadmConfig * max(1, len(alt_cfgs)))(*alt_cfgs)
        err = C.c_int(0)
        p = self.lib.orc_pipeline_create(sample_rate, len(ch), fft_size, int(use_denoiser), C.byref(main_cfg), alts,
                                         len(alt_cfgs), int(faithful), C.byref(err))
        if not p:
            raise ValueError(f"orc_pipeline_create -> {err.value}")
        try:
            if trace:
                self.lib.orc_pipeline_enable_trace(p, 1, 1, 1)
            if spectra:
                self.lib.orc_pipeline_enable_spectra(p, 1)
            ptrs = (_fp * len(ch))()
            for off in range(0, n, push):
                m = min(push, n - off)
                for c in range(len(ch)):
                    ptrs[c] = C.cast(ch[c].ctypes.data + 4 * off, _fp)
                self.lib.orc_pipeline_push(p, ptrs, m)
            res = {"segments": []}
            for k in range(1 + len(alt_cfgs)):
                cnt = self.lib.orc_pipeline_segments(p, k, None, 0)
                buf = np.zeros(cnt, SEG_DTYPE)
                self.lib.orc_pipeline_segments(p, k, buf.ctypes.data, cnt)
                res["segments"].append(buf)
            ne = self.lib.orc_pipeline_events(p, None, 0)
            ev = (CaptureEvent * max(1, ne))()
            self.lib.orc_pipeline_events(p, ev, ne)
            res["events"] = [(e.at_index, e.kind, e.sample_number) for e in ev[:ne]]
            if trace:
                cnt = self.lib.orc_pipeline_stft_records(p, None, 0)
                rec = np.zeros(cnt, REC_DTYPE)
                self.lib.orc_pipeline_stft_records(p, rec.ctypes.data, cnt)
                res["records"] = rec
                den = []
                for c in range(len(ch)):
                    m = self.lib.orc_pipeline_denoised(p, c, None, 0)
                    d = np.empty(m, np.float32)
                    self.lib.orc_pipeline_denoised(p, c, _f(d), m)
                    den.append(d)
                res["denoised"] = den
                m = self.lib.orc_pipeline_rnn_vad(p, None, 0)
                v = np.empty(m, np.float32)
                self.lib.orc_pipeline_rnn_vad(p, _f(v), m)
                res["rnn_vad"] = v
            if spectra:
                sp = []
                for c in range(len(ch)):
                    m = self.lib.orc_pipeline_spectra(p, c, None, 0)
                    d = np.empty(m, np.float32)
                    self.lib.orc_pipeline_spectra(p, c, _f(d), m)
                    sp.append(d.reshape(-1, fft_size // 2 + 1))
                res["spectra"] = sp
            return res
        finally:
            self.lib.orc_pipeline_destroy(p)

    def vadm_scan(self, cfg, records, sample_rate=48000, fft_size=2048):
        records = np.ascontiguousarray(records, REC_DTYPE)
        out = np.zeros(4096, SEG_DTYPE)
        n = self.lib.orc_vadm_scan(C.byref(cfg), sample_rate, fft_size, records.ctypes.data, records.size,
                                   out.ctypes.data, out.size)
        if n > out.size:
            out = np.zeros(n, SEG_DTYPE)
            self.lib.orc_vadm_scan(C.byref(cfg), sample_rate, fft_size, records.ctypes.data, records.size,
                                   out.ctypes.data, out.size)
        return out[:n].copy()


def have_reference():
    return os.path.exists(REF_SO)


def have_reference_fast():
    """The -O3 -march=x86-64-v3 build exists AND this host can run it (AVX2 + FMA)."""
    if not os.path.exists(REF_FAST_SO):
        return False
    try:
        flags = next(l for l in open("/proc/cpuinfo") if l.startswith("flags")).split()
    except (OSError, StopIteration):
        return False
    return all(f in flags for f in ("avx2", "fma", "bmi2"))


_synth = None


def log10_ex(ex):
    """(float)log10(1e-2 + ex) with the host libm, as lib/rnnoise/src/denoise.c:349 evaluates it."""
    synth(0, 0, 0, 0)                                   # loads the helper library
    ex = np.ascontiguousarray(ex, np.float32)
    out = np.empty_like(ex)
    _synth.orc_log10_ex.argtypes = [_fp, _fp, C.c_size_t]
    _synth.orc_log10_ex(_f(ex), _f(out), ex.size)
    return out


def synth(seed, stream_id, first_sample, n):
    """The seeded input generator (include/fvad_synth_core.h) on the host, without libfvad_cuda.so."""
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_SO):
            build("port")
        _synth = C.CDLL(SYNTH_SO)
        _synth.orc_synth_fill.argtypes = [_fp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_size_t]
    out = np.empty(n, np.float32)
    _synth.orc_synth_fill(_f(out), seed, stream_id, first_sample, n)
    return out
