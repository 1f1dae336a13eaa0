"""GPU vs the COMMITTED golden fixtures (tests/golden/*.npz, generated from the reference's own C build by
tools/make_golden.py): the CUDA path through the C ABI must reproduce the stored reference outputs directly, with no
oracle library in the loop -- this is what still holds on a box where oracle/_ref could not be built."""
import hashlib
import os

import numpy as np
import pytest

from conftest import synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def gold(name):
    return np.load(os.path.join(GOLD, name))


def _records_equal(got, want):
    return all(np.array_equal(got[k], want[k]) for k in ("index", "min_volume", "vol_ratio", "rnn_vad", "has_vad"))


def _segments_equal(got, want):
    return len(got) == len(want) and all(np.array_equal(got[k], want[k]) for k in ("sample_from", "sample_to", "rnn_vad", "vol_ratio"))


def _events(b, s):
    return np.array([(int(e["at_index"]), int(e["recording_state"]), int(e["sample_number"])) for e in b.events(s)], dtype=np.uint64).reshape(-1, 3)


def test_fft_golden(fv):
    z = gold("fft2048.npz")
    f = fv.FFT(2048, 48000)
    for i in range(len(z["x"])):
        assert np.array_equal(f.fft(z["x"][i], z["window"]), z["mags"][i])
        assert np.array_equal(f.fft(z["x"][i][:700], z["window"], second=z["x"][i][700:]), z["mags"][i])     # SplitSlice form
    f.deinit()


def test_denoiser_golden(fv):
    z = gold("denoiser_s21_6s.npz")
    x = synth(fv, int(z["stream"]), int(z["n"]) / 48000)
    assert np.array_equal(sha(x), z["input_sha"])
    out, vad = fv.DenoiserBatch(1).run(x[None, :])
    assert np.array_equal(vad[0], z["vad"]) and np.array_equal(sha(out[0]), z["out_sha"])
    assert np.array_equal(out[0][:z["out_head"].size], z["out_head"]) and np.array_equal(out[0][-z["out_tail"].size:], z["out_tail"])


def test_pipeline_golden_config1(fv):
    """BASELINE.json configs[0]: one synthetic 48 kHz mono stream, 60 s, default plan (+ one alt machine)."""
    z = gold("pipeline_s0_60s.npz")
    x = synth(fv, 0, 60)
    assert np.array_equal(sha(x), z["input_sha"])
    alt = fv.default_machine_config(speech_threshold_factor=9.0, min_vad_duration_sec=0.3, max_speech_gap_sec=1.0)
    flags = fv.TRACE_RECORDS | fv.TRACE_DENOISED | fv.TRACE_RNN_VAD
    for push in (48000, 240000, 1234 * 7):
        with fv.PipelineBatch(1, alt_vad_machine_configs=[alt], trace=flags) as b:
            for off in range(0, x.size, push):
                b.push_samples(x[None, off:off + push])
            assert np.array_equal(b.trace_rnn_vad(0), z["rnn_vad"])
            assert _records_equal(b.trace_records(0), z["records"])
            assert _segments_equal(b.segments(0, 0), z["segments0"]) and _segments_equal(b.segments(0, 1), z["segments1"])
            assert np.array_equal(_events(b, 0), z["events"])
            assert np.array_equal(sha(b.trace_denoised(0, 0)), z["denoised_sha"])


def test_pipeline_golden_stereo_shared_state(fv):
    z = gold("pipeline_stereo_s205_40s.npz")
    ch = np.stack([synth(fv, 205, 40), synth(fv, 206, 40) * z["gain1"]])
    assert np.array_equal(sha(ch), z["input_sha"])
    flags = fv.TRACE_RECORDS | fv.TRACE_DENOISED | fv.TRACE_RNN_VAD
    with fv.PipelineBatch(1, n_channels=2, trace=flags) as b:
        b.push_samples(ch[None, :, :])
        assert np.array_equal(b.trace_rnn_vad(0), z["rnn_vad"]) and _records_equal(b.trace_records(0), z["records"])
        assert _segments_equal(b.segments(0, 0), z["segments0"]) and np.array_equal(_events(b, 0), z["events"])
        assert np.array_equal(sha(np.stack([b.trace_denoised(0, 0), b.trace_denoised(0, 1)])), z["denoised_sha"])


def test_pipeline_golden_without_denoiser(fv):
    z = gold("pipeline_nodenoise_s107_40s.npz")
    x = synth(fv, 107, 40)
    assert np.array_equal(sha(x), z["input_sha"])
    cfg = fv.default_machine_config(speech_threshold_factor=4.0, initial_long_term_avg=0.05)
    with fv.PipelineBatch(1, use_denoiser=False, vad_machine_config=cfg, trace=fv.TRACE_RECORDS | fv.TRACE_SPECTRA) as b:
        b.push_samples(x[None, :])
        assert _records_equal(b.trace_records(0), z["records"]) and _segments_equal(b.segments(0, 0), z["segments0"])
        assert np.array_equal(_events(b, 0), z["events"])
        sp = b.trace_spectra(0, 0)
        assert np.array_equal(sha(sp), z["spectra_sha"]) and np.array_equal(sp[:3], z["spectra_head"])
