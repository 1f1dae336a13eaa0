"""The C-ABI boundary without a GPU: libfvad_cuda.so loads, exports every symbol include/fvad.h
declares, mirrors the reference's config defaults and error names, and FAILS LOUDLY (no CPU path)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fvad.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fvad_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(fv):
    names = declared_functions()
    assert len(names) >= 45, names
    out = subprocess.run(["nm", "-D", "--defined-only", fv.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\b(fvad_[a-z0-9_]+)\b", out))
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in include/fvad.h but not exported: {missing}"
    L = fv.lib()
    for n in names:                       # and the ctypes host side binds each of them
        assert hasattr(L, n), n


def test_product_does_not_link_the_oracle_or_emulation(fv):
    out = subprocess.run(["ldd", fv.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "rn_emu" not in out and "kiss" not in out and "rnnoise" not in out
    syms = subprocess.run(["nm", "-D", fv.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in syms and "rnnoise_process_frame" not in syms and "kiss_fft" not in syms
    # no product source mentions the oracle
    for dirpath, _, files in os.walk(os.path.join(ROOT, "formula-vad_b200")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".py", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in txt and "import oracle" not in txt and "from oracle" not in txt, f


def test_default_config_matches_reference(fv):
    c = fv.default_machine_config()       # VADMachine.zig:18-39
    assert (c.speech_min_freq, c.speech_max_freq, c.long_term_speech_avg_sec) == (100.0, 1500.0, 180.0)
    assert c.has_initial_long_term_avg == 1 and c.initial_long_term_avg == 0.005
    assert c.short_term_speech_avg_sec == np.float32(0.2) and c.speech_threshold_factor == 18.0
    assert c.channel_vol_ratio_avg_sec == 0.5 and c.channel_vol_ratio_threshold == 0.5
    assert c.min_consecutive_sec_to_open == np.float32(0.2) and c.max_speech_gap_sec == 2.0
    assert c.min_vad_duration_sec == np.float32(0.7)
    cfg = fv.Config()
    fv.lib().fvad_default_config(C.byref(cfg), 48000, 2)       # AudioPipeline.zig:20-26, VAD.zig:17-23
    assert cfg.sample_rate == 48000 and cfg.n_channels == 2 and cfg.vad_config.fft_size == 2048
    assert cfg.vad_config.use_denoiser == 1 and cfg.skip_processing == 0 and cfg.buffer_length == 0


def test_argument_errors_carry_reference_names(fv):
    with pytest.raises(fv.FvadError) as e:                       # VAD.zig:101-104
        fv.PipelineBatch(1, sample_rate=44100)
    assert e.value.name == "InvalidSampleRate"
    with pytest.raises(fv.FvadError) as e:                       # FFT.zig:29-31
        fv.PipelineBatch(1, fft_size=1023)
    assert e.value.name == "InvalidFFTSize"
    with pytest.raises(fv.FvadError) as e:
        fv.FFT(0, 48000)
    assert e.value.name == "InvalidFFTSize"
    assert fv.Denoiser.get_frame_size() == 480                   # Denoiser.zig:68-70
    L = fv.lib()
    names = {L.fvad_status_name(-i).decode() for i in range(1, 16)}
    assert {"InvalidSampleRate", "InvalidFFTSize", "InvalidSamplesLength", "InvalidWindowLength", "InvalidResultLength",
            "OutOfRange", "NegativeFrequency", "InvalidFrameSize", "InvalidResultSize", "NotInitialized",
            "ChannelCountMismatch", "OutOfMemory", "CudaError"} <= names


def test_no_cpu_fallback(fv):
    """Without a CUDA device every compute entry point must fail with CudaError, never compute on the host."""
    if fv.lib().fvad_device_count() > 0:
        pytest.skip("a GPU is present")
    for make in (lambda: fv.PipelineBatch(1), lambda: fv.PipelineBatch(2, use_denoiser=False), lambda: fv.FFT(2048, 48000),
                 fv.Denoiser, lambda: fv.DenoiserBatch(3)):
        with pytest.raises(fv.FvadError) as e:
            make()
        assert e.value.name == "CudaError", e.value
    rec = np.zeros((1, 4), fv.REC_DTYPE)
    with pytest.raises(fv.FvadError) as e:
        fv.vadm_sweep(rec, [fv.default_machine_config()])
    assert e.value.name == "CudaError"


def test_host_side_tables_and_synth(fv, port):
    """Host helpers that are legitimately CPU code: window tables and the seeded generator."""
    assert np.array_equal(fv.hann_window_periodic(2048), port.hann(2048))
    a = fv.synth_host(0xF1AD5EED00000000, 5, 0, 4800)
    b = fv.synth_host(0xF1AD5EED00000000, 5, 2400, 2400)
    assert np.array_equal(a[2400:], b) and np.abs(a).max() <= 1.0      # pure function of (seed, stream, n)
    assert not np.array_equal(a, fv.synth_host(0xF1AD5EED00000000, 6, 0, 4800))


def test_sass_exactness_guard():
    """The built library's STFT kernels use packed f32x2 arithmetic + TMA bulk copies and contain no contracted
    multiply-add (every FFMA2 is fma(a, b, -0), see stft.cu prod2): tools/check_sass.sh inspects the SASS."""
    import os
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if shutil.which("cuobjdump") is None and not os.path.exists("/usr/local/cuda/bin/cuobjdump"):
        pytest.skip("cuobjdump not available")
    env = dict(os.environ, PATH=os.environ.get("PATH", "") + ":/usr/local/cuda/bin")
    r = subprocess.run(["bash", os.path.join(root, "tools", "check_sass.sh")], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr


def test_spectrogram_shape_host_logic(fv):
    """Spectrogram.compute's bookkeeping (Spectrogram.zig:34-39,66,81) needs no device."""
    i = fv.Spectrogram.shape(48000 * 3, 48000, 2048, 1024)
    assert (i.height, i.width) == (1025, (48000 * 3 - 1024) // 1024)
    assert i.bin_width == np.float32(23.4375) and i.nyquist_freq == 24000.0
    assert i.length_sec == ((i.width - 1) * 1024 + 2048) / 48000
    for args, name in (((100, 48000, 2048, 2048), "InsufficientSamples"), ((4096, 48000, 2048, 512), "OutOfRange"),
                       ((4096, 48000, 2047, 512), "InvalidFFTSize")):
        with pytest.raises(fv.FvadError) as e:
            fv.Spectrogram.shape(*args)
        assert e.value.name == name
