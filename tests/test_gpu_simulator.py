"""GPU: the simulator-shaped driver (fvad/simulator.py) on a plan whose streams differ in length and
format.  Every instance's segment list must equal the oracle pipeline run on that stream ALONE and UNPADDED (the driver
pads shorter streams inside one batch and removes the padding's effect), the Audacity text must equal the oracle
restatement's, and the statistics must equal the oracle's bit for bit."""
import json
import os

import numpy as np
import pytest

import evaluator_ref as E
from conftest import synth, assert_segments_equal
from test_simulator_host import _write_pcm

pytestmark = pytest.mark.gpu


def test_plan_with_ragged_streams(fv, checker, tmp_path):
    from fvad import simulator as sim
    from fvad import evaluator as ev
    lengths = [48000 * 50, 48000 * 44 + 1234, 48000 * 31 + 77, 48000 * 38]
    insts = []
    rng = np.random.default_rng(5)
    for i, n in enumerate(lengths):
        x = synth(fv, 300 + i, n / 48000 + 0.01)[:n]
        if i == 1:
            _write_pcm(tmp_path / f"s{i}.wav", x[None, :], 16)                       # one 16-bit file: decoded samples differ from x
        else:
            sim.write_wav_f32(str(tmp_path / f"s{i}.wav"), x[None, :])
        t, lab = 0.0, []
        for _ in range(4):
            t += rng.uniform(2, 9); d = rng.uniform(1, 6); lab.append((t, t + d)); t += d
        (tmp_path / f"s{i}.txt").write_text("".join(f"{a:.3f}\t{b:.3f}\tref\n" for a, b in lab))
        insts.append({"name": f"car{i}", "audio_path": f"s{i}.wav", "ref_path": f"s{i}.txt"})
    plan = {"instances": insts, "config": {"vad_config": {"vad_machine_config": {"min_vad_duration_sec": 0.5}}, "output_dir": "sim-out"}}
    (tmp_path / "plan.json").write_text(json.dumps(plan))

    s = sim.Simulation(str(tmp_path / "plan.json"))
    s.run_all(chunk_seconds=17)                                                       # several ragged chunks
    stats, texts, report = s.results()
    assert s.save_output(texts, report)

    cfg = checker.default_config(min_vad_duration_sec=0.5)
    stat_cfg = dict(ignore_shorter_than_sec=np.float32(0.5), extrude_start=5, extrude_end=10, fill_gaps=5)   # simulator.zig:121-126
    total = dropped = 0
    for i, inst in enumerate(s.instances):
        pcm = inst.audio.read(0, inst.audio.n_frames)
        assert pcm.shape == (1, lengths[i])
        want = checker.run_pipeline([pcm[0]], main_cfg=cfg)["segments"][0]
        assert_segments_equal(inst.vad_segments, want)
        assert np.array_equal(inst.vad_segments["rnn_vad"], want["rnn_vad"])
        total += len(want)
        ins, refs = E.init_and_run(E.vad_segments_to_speech(want, 48000), E.parse_audacity((tmp_path / f"s{i}.txt").read_text()))
        assert texts[inst.name] == E.serialize_audacity(ins, refs)
        wstats = E.from_evaluator(ins, refs, stat_cfg)
        for k in ev.STATS_DTYPE.names:
            g, w = stats[i][k], np.float32(wstats[k])
            assert (np.isnan(g) and np.isnan(w)) or g.tobytes() == w.tobytes(), (i, k, g, w)
        assert open(os.path.join(s.out_path, f"{inst.name}-audacity.txt")).read() == texts[inst.name]
        # recordings (Recorder.zig): one clip per started -> completed pair whose end the stream reaches
        oev = checker.run_pipeline([pcm[0]], main_cfg=cfg, trace=True)["events"]
        want_clips, start = [], None
        for (_, state, sample) in oev:
            if state == 1: start = sample
            elif state == 3: start = None
            elif state == 2 and start is not None:
                if sample <= lengths[i]: want_clips.append((start, sample))
                else: dropped += 1
                start = None
        clips = list(inst.recordings())
        assert [(c["start"], c["end"]) for c in clips] == want_clips
        for c in clips:
            assert c["channel"] == 0 and np.array_equal(c["pcm"], pcm[0, c["start"]:c["end"]])
        assert len(os.listdir(os.path.join(s.out_path, inst.name))) == len(clips)
    assert total > 0
    assert "=> Performance Report" in report and "car3" in report and os.path.isfile(os.path.join(s.out_path, "report.txt"))
