"""GPU parity for BASELINE config 5 (Evaluator VAD-threshold sweep): many VADMachine configurations over SHARED
features (VAD.zig:171-183,375-380).  fvad_vadm_sweep over the pipeline's own STFT records must give, for every
(stream, config), the segment list the oracle's alt machines produce inside a full pipeline run; the sweep's consumer
(fvad_eval_stats_batch) is then run on the sweep's output layout."""
import itertools

import numpy as np
import pytest

import evaluator_ref as E
from conftest import synth, assert_segments_equal

pytestmark = pytest.mark.gpu


def test_sweep_matches_alt_machines(fv, checker):
    from fvad import evaluator as ev
    S, secs = 3, 70
    grid = list(itertools.product((6.0, 12.0, 18.0, 30.0), (0.1, 0.2, 0.4), (0.1, 0.2), (0.5, 2.0), (0.3, 0.7)))   # 96 configs
    keys = ("speech_threshold_factor", "short_term_speech_avg_sec", "min_consecutive_sec_to_open", "max_speech_gap_sec", "min_vad_duration_sec")
    gcfgs = [fv.default_machine_config(**dict(zip(keys, g))) for g in grid]
    ocfgs = [checker.default_config(**dict(zip(keys, g))) for g in grid]
    pcm = np.stack([synth(fv, 400 + s, secs) for s in range(S)])
    with fv.PipelineBatch(S, trace=fv.TRACE_RECORDS) as b:
        b.push_samples(pcm)
        b.sync()
        recs = np.stack([b.trace_records(s) for s in range(S)])
    got = fv.vadm_sweep(recs, gcfgs)
    total = distinct = 0
    for s in range(S):
        want = checker.run_pipeline([pcm[s]], alt_cfgs=ocfgs)["segments"]
        seen = set()
        for k in range(len(grid)):
            assert_segments_equal(got[s][k], want[1 + k])
            assert np.array_equal(got[s][k]["rnn_vad"], want[1 + k]["rnn_vad"]) and np.array_equal(got[s][k]["vol_ratio"], want[1 + k]["vol_ratio"])
            total += len(want[1 + k])
            seen.add(tuple(want[1 + k]["sample_from"]) + tuple(want[1 + k]["sample_to"]))
        distinct += len(seen)
    assert total > 0 and distinct > S, "the grid must change the decisions"

    # the sweep's consumer on the same layout: statistics of every (stream, config) against per-stream labels
    refs = [np.array([(5.0, 9.0), (20.0, 26.5), (40.0, 41.0)], dtype=ev.REF_DTYPE) for _ in range(S)]
    per = [got[s][k] for s in range(S) for k in range(len(grid))]
    segs, counts = ev.stack_segments(per)
    set_ref = np.repeat(np.arange(S, dtype=np.uint32), len(grid))
    cfg = dict(ignore_shorter_than_sec=0.7, extrude_start=5, extrude_end=10, fill_gaps=5)
    st = ev.stats_batch(segs, counts, set_ref, refs, 48000, ev.StatConfig(0.7, 5, 10, 5))
    for i in range(0, len(per), 7):
        ins, rf = E.init_and_run(E.vad_segments_to_speech(per[i], 48000), [E.Seg(r["from_sec"], r["to_sec"], "ref") for r in refs[set_ref[i]]])
        w = E.from_evaluator(ins, rf, cfg)
        for k in ev.STATS_DTYPE.names:
            g, x = st[i][k], np.float32(w[k])
            assert (np.isnan(g) and np.isnan(x)) or g.tobytes() == x.tobytes(), (i, k, g, x)
