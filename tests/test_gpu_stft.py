"""GPU parity: FFT.fft / batched STFT + speech-band sum and the use_denoiser=false pipeline
against the oracle (bit-exact spectra, band sums, segment boundaries)."""
import numpy as np
import pytest

from conftest import synth, assert_segments_equal

pytestmark = pytest.mark.gpu


def test_fft_single_bit_exact(fv, checker):
    rng = np.random.default_rng(7)
    f = fv.FFT(2048, 48000)
    w = fv.hann_window_periodic(2048)
    assert np.array_equal(w, checker.hann(2048))
    for k in range(4):
        x = (rng.standard_normal(2048) * 10.0 ** (-k)).astype(np.float32)
        got = f.fft(x[:700], w, second=x[700:])          # SplitSlice halves
        want = checker.fft(x, window=w)
        assert np.array_equal(got, want), np.abs(got - want).max()
    # a non-Hann window goes through the same path (FFT.fft takes the window as an argument)
    w2 = np.ones(2048, np.float32)
    x = rng.standard_normal(2048).astype(np.float32)
    assert np.array_equal(f.fft(x, w2), checker.fft(x, window=w2))


def test_fft_errors(fv):
    with pytest.raises(fv.FvadError) as e:
        fv.FFT(0, 48000)
    assert e.value.name == "InvalidFFTSize"
    with pytest.raises(fv.FvadError) as e:
        fv.FFT(1023, 48000)
    assert e.value.name == "InvalidFFTSize"
    f = fv.FFT(2048, 48000)
    w = fv.hann_window_periodic(2048)
    with pytest.raises(fv.FvadError) as e:
        f.fft(np.zeros(2047, np.float32), w)
    assert e.value.name == "InvalidSamplesLength"
    with pytest.raises(fv.FvadError) as e:
        f.fft(np.zeros(2048, np.float32), w[:100])
    assert e.value.name == "InvalidWindowLength"
    with pytest.raises(fv.FvadError) as e:
        f.fft(np.zeros(2048, np.float32), w, result_len=1024)
    assert e.value.name == "InvalidResultLength"
    assert f.bin_count() == 1025 and f.freq_to_bin(100.0) == 4 and f.freq_to_bin(1500.0) == 64
    with pytest.raises(fv.FvadError) as e:
        f.freq_to_bin(30000.0)
    assert e.value.name == "OutOfRange"
    with pytest.raises(fv.FvadError) as e:
        f.freq_to_bin(-1.0)
    assert e.value.name == "NegativeFrequency"


def test_stft_batch_band_and_spectra(fv, checker):
    import torch
    rows, frames = 5, 9
    x = np.stack([synth(fv, 10 + r, frames * 2048 / 48000) for r in range(rows)])
    d = torch.from_numpy(x).cuda()
    band = torch.zeros(rows, frames, device="cuda")
    spec = torch.zeros(rows, frames, 1025, device="cuda")
    f = fv.FFT(2048, 48000)
    f.stft_batch(d.data_ptr(), rows, x.shape[1], frames, 2048, 4, 64, band.data_ptr(), spec.data_ptr())
    band2 = torch.zeros(rows, frames, device="cuda")
    f.stft_batch(d.data_ptr(), rows, x.shape[1], frames, 2048, 4, 64, band2.data_ptr())      # pruned kernel
    torch.cuda.synchronize()
    w = checker.hann(2048)
    for r in range(rows):
        for j in range(frames):
            want = checker.fft(x[r, j * 2048:(j + 1) * 2048], window=w)
            assert np.array_equal(spec[r, j].cpu().numpy(), want)
            s = np.float32(0)
            for i in range(4, 65):
                s = np.float32(s + want[i])
            assert band[r, j].item() == s and band2[r, j].item() == s


@pytest.mark.parametrize("n_channels", [1, 2])
def test_pipeline_without_denoiser(fv, checker, n_channels):
    S, secs = 3, 40
    pcm = np.stack([np.stack([synth(fv, 100 + 7 * s + c, secs) * (1.0 if c == 0 else 0.8) for c in range(n_channels)])
                    for s in range(S)])
    cfg = fv.default_machine_config(speech_threshold_factor=4.0, initial_long_term_avg=0.05)
    alt = fv.default_machine_config(speech_threshold_factor=3.0, speech_min_freq=200.0, speech_max_freq=3000.0,
                                    min_vad_duration_sec=0.3, initial_long_term_avg=None,
                                    long_term_speech_avg_sec=20.0, max_speech_gap_sec=0.5)
    with fv.PipelineBatch(S, n_channels=n_channels, use_denoiser=False, vad_machine_config=cfg,
                          alt_vad_machine_configs=[alt], trace=fv.TRACE_RECORDS) as b:
        # ragged pushes exercise the carry between tiles
        off = 0
        for n in (1000, 48000, 5, 2047, 2049, 48000 * 3):
            b.push_samples(pcm[:, :, off:off + n]); off += n
        b.push_samples(pcm[:, :, off:])
        from oracle import VadmConfig
        ocfg = checker.default_config(speech_threshold_factor=4.0, initial_long_term_avg=0.05)
        oalt = checker.default_config(speech_threshold_factor=3.0, speech_min_freq=200.0, speech_max_freq=3000.0,
                                      min_vad_duration_sec=0.3, initial_long_term_avg=None,
                                      long_term_speech_avg_sec=20.0, max_speech_gap_sec=0.5)
        total = 0
        for s in range(S):
            want = checker.run_pipeline(list(pcm[s]), use_denoiser=False, main_cfg=ocfg, alt_cfgs=[oalt], trace=True)
            rec = b.trace_records(s)
            assert np.array_equal(rec["index"], want["records"]["index"])
            assert np.array_equal(rec["min_volume"], want["records"]["min_volume"])
            assert np.array_equal(rec["vol_ratio"], want["records"]["vol_ratio"])
            for k in range(2):
                assert_segments_equal(b.segments(s, k), want["segments"][k])
                total += len(want["segments"][k])
            ev = b.events(s)
            assert [(int(e["at_index"]), int(e["recording_state"]), int(e["sample_number"])) for e in ev] == want["events"]
        assert total > 0, "test input produced no segments: not a meaningful check"


@pytest.mark.parametrize("n_fft", [4, 30, 256, 480, 960, 1024, 1920, 4096, 8192, 16384])
def test_fft_other_sizes_bit_exact(fv, checker, n_fft):
    """FFT.init accepts any even n_fft (src/FFT.zig:28-58).  Sizes other than 2048 run KissFFT's generic plan
    (csrc/stft_generic.cu: radix 4 / 2 / 3 / 5 stages in kf_factor's order, kiss_fftr's split): magnitudes and the batched
    band sums must equal the oracle's FFT.fft bit for bit, with the size's Hann window and with a non-Hann one."""
    import torch
    if checker.kind == "port" and (n_fft & (n_fft - 1)):
        pytest.skip("the plain-C port of the oracle covers power-of-two sizes; the reference build covers all")
    rng = np.random.default_rng(n_fft)
    f = fv.FFT(n_fft, 48000)
    assert f.bin_count() == n_fft // 2 + 1
    w = fv.hann_window_periodic(n_fft)
    assert np.array_equal(w, checker.hann(n_fft))
    for k, win in enumerate((w, (0.25 + 0.5 * rng.random(n_fft)).astype(np.float32))):
        x = (rng.standard_normal(n_fft) * 10.0 ** (-k)).astype(np.float32)
        got = f.fft(x[:n_fft // 3], win, second=x[n_fft // 3:])
        want = checker.fft(x, win, n_fft=n_fft)
        assert np.array_equal(got, want), (n_fft, k, np.abs(got - want).max())
    # batched: rows x overlapping frames, band sum + spectra
    rows, frames, hop = 3, 5, max(2, n_fft // 2)
    n = hop * (frames - 1) + n_fft
    pcm = (rng.standard_normal((rows, n)) * 0.1).astype(np.float32)
    d = torch.from_numpy(pcm).cuda()
    lo, hi = (1, n_fft // 4) if n_fft >= 8 else (0, 1)
    band = torch.zeros(rows, frames, dtype=torch.float32, device="cuda")
    spec = torch.zeros(rows, frames, n_fft // 2 + 1, dtype=torch.float32, device="cuda")
    f.stft_batch(d.data_ptr(), rows, n, frames, hop, lo, hi, band.data_ptr(), spec.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for r in range(rows):
        for j in range(frames):
            want = checker.fft(pcm[r, j * hop:j * hop + n_fft], w, n_fft=n_fft)
            assert np.array_equal(spec[r, j].cpu().numpy(), want)
            acc = np.float32(0)
            for v in want[lo:hi + 1]:
                acc = np.float32(acc + v)
            assert np.float32(band[r, j].item()) == acc
    f.deinit()


def test_fft_unsupported_sizes_say_so(fv):
    with pytest.raises(fv.FvadError) as e:                       # complex length 7: kf_bfly_generic is not implemented
        fv.FFT(14, 48000)
    assert e.value.name == "Unsupported" and "prime factor" in str(e.value)
    with pytest.raises(fv.FvadError) as e:
        fv.FFT(1023, 48000)
    assert e.value.name == "InvalidFFTSize"


@pytest.mark.parametrize("fft_size,use_denoiser", [(1024, True), (4096, True), (960, True), (512, False), (4096, False)])
def test_pipeline_other_fft_sizes(fv, checker, fft_size, use_denoiser):
    """VAD.Config.fft_size other than 2048 through the whole pipeline (re-framing 480 -> fft_size, band bins of the size,
    window lengths in STFT frames of the size): records, segments and events equal the oracle's."""
    from conftest import synth
    if checker.kind == "port" and (fft_size & (fft_size - 1)):
        pytest.skip("the plain-C port of the oracle covers power-of-two sizes; the reference build covers all")
    S, secs = 2, 40
    pcm = np.stack([synth(fv, 900 + s, secs) for s in range(S)])
    kw = dict(speech_threshold_factor=6.0, min_vad_duration_sec=0.3) if not use_denoiser else {}
    if not use_denoiser:
        kw["initial_long_term_avg"] = 0.05
    total = 0
    with fv.PipelineBatch(S, fft_size=fft_size, use_denoiser=use_denoiser, vad_machine_config=fv.default_machine_config(**kw),
                          trace=fv.TRACE_RECORDS) as b:
        b.push_samples(pcm[:, :48000 * 7 + 11])
        b.push_samples(pcm[:, 48000 * 7 + 11:])
        for s in range(S):
            want = checker.run_pipeline([pcm[s]], use_denoiser=use_denoiser, fft_size=fft_size, main_cfg=checker.default_config(**kw), trace=True)
            rec = b.trace_records(s)
            for k in ("index", "min_volume", "vol_ratio", "rnn_vad"):
                assert np.array_equal(rec[k], want["records"][k]), (fft_size, k)
            assert b.segments(s, 0).tobytes() == want["segments"][0].tobytes()
            ev = b.events(s)
            assert [(int(e["at_index"]), int(e["recording_state"]), int(e["sample_number"])) for e in ev] == want["events"]
            total += len(want["segments"][0])
    assert total > 0
