"""Host logic of the simulator-shaped driver (fvad/simulator.py; reference src/simulator.zig, SimulationInstance.zig):
WAV decoding as sf_readf_float normalises, plan parsing, and the bookkeeping that removes a zero-padded tail."""
import json
import os
import struct

import numpy as np
import pytest


def _write_pcm(path, x, bits, rate=48000):
    a = np.atleast_2d(x)
    ch, n = a.shape
    if bits == 16:
        q = np.clip(np.round(a * 32768), -32768, 32767).astype("<i2")
        data, want = q.T.tobytes(), q.astype(np.float32) / np.float32(32768)
    elif bits == 24:
        q = np.clip(np.round(a * 8388608), -8388608, 8388607).astype(np.int32)
        b = np.stack([(q.T & 0xFF), ((q.T >> 8) & 0xFF), ((q.T >> 16) & 0xFF)], -1).astype(np.uint8)
        data, want = b.tobytes(), q.astype(np.float32) / np.float32(8388608)
    else:
        q = (np.clip(np.round(a * 128), -128, 127) + 128).astype(np.uint8)
        data, want = q.T.tobytes(), (q.astype(np.float32) - 128) / np.float32(128)
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(data) + 12) + b"WAVE")
        f.write(b"LIST" + struct.pack("<I", 4) + b"INFO")                       # an unknown chunk before fmt
        f.write(b"fmt " + struct.pack("<IHHIIHH", 16, 1, ch, rate, rate * ch * bits // 8, ch * bits // 8, bits))
        f.write(b"data" + struct.pack("<I", len(data)) + data)
    return want


@pytest.mark.parametrize("bits", [8, 16, 24])
def test_wav_pcm_decoding(fv, tmp_path, bits):
    from fvad import simulator as sim
    rng = np.random.default_rng(bits)
    x = rng.uniform(-1, 1, (2, 5000)).astype(np.float32)
    want = _write_pcm(tmp_path / "a.wav", x, bits)
    w = sim.WavFile(str(tmp_path / "a.wav"))
    assert (w.n_channels, w.sample_rate, w.n_frames) == (2, 48000, 5000)
    assert np.array_equal(w.read(0, 5000), want)
    assert np.array_equal(w.read(4990, 100), want[:, 4990:])                      # short read at the end
    assert w.read(5000, 10).shape == (2, 0)


def test_wav_float_roundtrip_and_errors(fv, tmp_path):
    from fvad import simulator as sim
    x = np.random.default_rng(1).uniform(-1, 1, (1, 777)).astype(np.float32)
    sim.write_wav_f32(str(tmp_path / "f.wav"), x)
    assert np.array_equal(sim.WavFile(str(tmp_path / "f.wav")).read(0, 777), x)
    (tmp_path / "x.ogg").write_bytes(b"OggS" + b"\0" * 64)
    with pytest.raises(sim.UnsupportedAudioFormat):
        sim.WavFile(str(tmp_path / "x.ogg"))


def test_plan_parsing_and_tail_bookkeeping(fv, tmp_path):
    from fvad import simulator as sim
    sim.write_wav_f32(str(tmp_path / "a.wav"), np.zeros((1, 48000), np.float32))
    (tmp_path / "a.txt").write_text("3.5\t4.25\tlabel\n1.0\t2.0\n")
    plan = {"instances": [{"name": "A", "audio_path": "a.wav", "ref_path": "a.txt"}],
            "config": {"vad_config": {"vad_machine_config": {"speech_threshold_factor": 9.0, "initial_long_term_avg": None, "unknown": 1}},
                       "output_dir": "out", "audio_read_frame_count": 48000, "unknown_key": True}}
    (tmp_path / "plan.json").write_text(json.dumps(plan))
    s = sim.Simulation(str(tmp_path / "plan.json"))
    assert s.machine_config.speech_threshold_factor == 9.0 and s.machine_config.has_initial_long_term_avg == 0
    assert s.fft_size == 2048 and s.use_denoiser is True
    assert list(s.instances[0].reference["from_sec"]) == [1.0, 3.5]                 # ordered by start (Evaluator.zig:108-109)
    assert os.path.isfile(os.path.join(s.out_path, "plan.json"))
    # frames the reference reaches: whole 480-steps, then whole 2048-frames
    assert sim.frames_decided(48000) == 23 and sim.frames_decided(6239) == 2 and sim.frames_decided(6240) == 3
    seg = np.zeros(3, fv.SEG_DTYPE)
    seg["sample_to"] = [10, 20, 30]
    evs = np.zeros(5, fv.EVENT_DTYPE)
    evs["recording_state"] = [1, 2, 2, 1, 2]
    evs["at_index"] = [0, 2048 * 5, 2048 * 22, 2048 * 22, 2048 * 23]
    kept = sim.keep_decided_inside(seg, evs, 48000)
    assert list(kept["sample_to"]) == [10, 20]                                      # frame 23 lies in the padding


def test_rms_volume_is_the_in_order_f32_sum(fv):
    from fvad import simulator as sim
    x = np.random.default_rng(2).uniform(-1, 1, 4801).astype(np.float32)
    s = np.float32(0)
    for v in x:
        s = np.float32(s + np.float32(v * v))                                       # audio_utils.zig:16-21
    assert sim.rms_volume(x) == np.float32(np.sqrt(np.float32(s / np.float32(x.size))))
