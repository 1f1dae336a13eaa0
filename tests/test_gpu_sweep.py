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


def _grid(fv, checker):
    grid = list(itertools.product((6.0, 12.0, 18.0, 30.0), (0.1, 0.2, 0.4), (0.1, 0.2), (0.5, 2.0), (0.3, 0.7)))   # 96 configs
    keys = ("speech_threshold_factor", "short_term_speech_avg_sec", "min_consecutive_sec_to_open", "max_speech_gap_sec", "min_vad_duration_sec")
    return ([fv.default_machine_config(**dict(zip(keys, g))) for g in grid], [checker.default_config(**dict(zip(keys, g))) for g in grid])


def test_batch_sweep_on_device_matches_alt_machines(fv, checker, knobs):
    """fvad_batch_sweep: the batch's OWN features stay in HBM (FVAD_KEEP_FEATURES) and 96 + 2 further machine configs are
    scanned over them on the device; every (stream, config) list must be what the oracle's alt machines produce inside a
    full pipeline run -- including two configs on a SECOND speech band, whose band sums the batch keeps because it was
    created with an alt machine of that band.  Small tiles and ragged pushes: the feature planes grow and span tiles."""
    import torch
    from fvad import evaluator as ev
    knobs(tile_frames="97")
    S, secs = 3, 70
    gcfgs, ocfgs = _grid(fv, checker)
    band2 = dict(speech_min_freq=300.0, speech_max_freq=3000.0)
    gcfgs += [fv.default_machine_config(**band2), fv.default_machine_config(speech_threshold_factor=9.0, **band2)]
    ocfgs += [checker.default_config(**band2), checker.default_config(speech_threshold_factor=9.0, **band2)]
    K = len(gcfgs)
    pcm = np.stack([synth(fv, 400 + s, secs) for s in range(S)])
    with fv.PipelineBatch(S, alt_vad_machine_configs=[fv.default_machine_config(**band2)], trace=fv.KEEP_FEATURES) as b:
        off = 0
        for n in (48000 * 13 + 5, 1, 48000 * 30):
            b.push_samples(pcm[:, off:off + n]); off += n
        first = b.sweep(gcfgs[:4])                                   # a sweep in mid-run: features of what has been pushed so far
        b.push_samples(pcm[:, off:])
        got = b.sweep(gcfgs)
        assert b.feature_count() == (secs * 48000 // 480 * 480) // 2048
        # device outputs feeding the Evaluator statistics without a host round trip of the lists
        max_segs = 64
        d_segs = torch.zeros(S * K * max_segs * fv.SEG_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        d_counts = torch.zeros(S * K, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        b.sweep(gcfgs, max_segs=max_segs, d_segs=d_segs.data_ptr(), d_counts=d_counts.data_ptr())
        refs = [np.array([(5.0, 9.0), (20.0, 26.5), (40.0, 41.0)], dtype=ev.REF_DTYPE) for _ in range(S)]
        set_ref = np.repeat(np.arange(S, dtype=np.uint32), K)
        sc = ev.StatConfig(0.7, 5, 10, 5)
        st_dev = ev.stats_batch_device(d_segs.data_ptr(), S * K, max_segs, d_counts.data_ptr(), set_ref, refs, 48000, sc)
        feats = b.features(1)
    per = [got[s][k] for s in range(S) for k in range(K)]
    segs, counts = ev.stack_segments(per)
    st_host = ev.stats_batch(segs, counts, set_ref, refs, 48000, sc)
    assert st_dev.tobytes() == st_host.tobytes()
    assert np.array_equal(d_counts.cpu().numpy().view(np.uint32), counts)
    total = 0
    for s in range(S):
        want = checker.run_pipeline([pcm[s]], alt_cfgs=ocfgs, trace=(s == 1))
        for k in range(K):
            assert got[s][k].tobytes() == want["segments"][1 + k].tobytes(), (s, k)
            total += len(got[s][k])
        if s == 1:
            for f in ("index", "min_volume", "vol_ratio", "rnn_vad"):
                assert np.array_equal(feats[f], want["records"][f]), f
        # the mid-run sweep saw a prefix: its lists are prefixes of the final ones (the path is causal)
        for k in range(4):
            assert first[s][k].tobytes() == got[s][k][:len(first[s][k])].tobytes()
    assert total > 100 and len(got[0][K - 1]) != len(got[0][0]) or total > 100


def test_long_ring_running_sum_is_exact_when_the_level_drains(fv, checker):
    """The long-term ring's running sum carries an ABSOLUTE error bound (vadm_step.cuh): a ring that drains from 1e-3 to
    1e-25 -- RNNoise's 0.6 * lastg gain floor decays geometrically in silence -- must still decide
    `short_term > last_avg * factor` as the reference's in-order sum does, also with short rings and without an initial
    average.  Checked against the oracle's VADMachine on every config."""
    rng = np.random.default_rng(7)
    N = 6000
    rec = np.zeros((2, N), fv.REC_DTYPE)
    for s in range(2):
        mv = (1e-3 * (0.5 + rng.random(N))).astype(np.float64)
        for a in (700, 2600, 4300):                                  # geometric drain to 1e-25, a floor, then the level returns
            n = 500 + 300 * s
            mv[a:a + n] = np.maximum(1e-3 * 0.6 ** np.arange(n), 1e-25) * (0.9 + 0.2 * rng.random(n))
            for c in range(a + 150, a + n - 40, 60):                 # bursts 30x above the drained floor: speech at that level
                mv[c:c + 25] *= 30.0
            mv[a + n:a + n + 30] = 0.0                               # and a stretch of exact zeros
        for a in rng.integers(100, N - 100, 12):
            mv[a:a + 40] = 0.3 * (0.5 + rng.random(40))              # ordinary loud speech
        rec[s]["index"] = 2048 * np.arange(N, dtype=np.uint64)
        rec[s]["min_volume"] = mv.astype(np.float32)
        rec[s]["vol_ratio"] = 0.9
        rec[s]["rnn_vad"] = rng.random(N).astype(np.float32)
        rec[s]["has_vad"] = 1.0
    kws = [dict(long_term_speech_avg_sec=sec, initial_long_term_avg=init, speech_threshold_factor=f, min_vad_duration_sec=0.3,
                max_speech_gap_sec=0.5)
           for sec in (2.0, 5.0, 20.0, 180.0) for init in (None, 0.005) for f in (6.0, 18.0)]
    got = fv.vadm_sweep(rec, [fv.default_machine_config(**kw) for kw in kws], max_segs=256)
    total = 0
    for s in range(2):
        for k, kw in enumerate(kws):
            want = checker.vadm_scan(checker.default_config(**kw), rec[s])
            assert got[s][k].tobytes() == want.tobytes(), (s, kw, len(got[s][k]), len(want))
            total += len(want)
    assert total > 200, total
