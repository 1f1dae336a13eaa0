"""Evaluator.  CPU part: the oracle restatement (oracle/evaluator_ref.py) replays the reference's
own unit tests (src/Evaluator/statistics.zig:286-360) and the product's host code (formats, report numbers) is checked
against it.  GPU part: fvad_eval_stats_batch equals the oracle bit for bit on random result sets."""
import numpy as np
import pytest

import evaluator_ref as E


def _cfg(**kw):
    c = dict(ignore_shorter_than_sec=0.0, extrude_start=0.0, extrude_end=0.0, fill_gaps=0.0)
    c.update(kw)
    return c


def test_oracle_matches_reference_unit_tests():
    """statistics.zig `calcFalsePositiveSec #1` (:286-322) and `#2` (:324-360): expectApproxEqAbs(fp, 0 | 3, 0.001)."""
    for vad_to, want in ((6, 0.0), (10, 3.0)):
        refs = [E.Seg(2, 3, "ref"), E.Seg(4, 5, "ref")]
        vad = E.Seg(1, vad_to, "vad")
        vad.opposite = refs
        fp = E.calc_false_positive_sec(vad, _cfg(extrude_start=2, extrude_end=2, fill_gaps=2))
        assert abs(float(fp) - want) <= 0.001


def test_zig_float_formatting():
    assert E.zig_fixed(np.float32(1.5), 4) == "1.5000" and E.zig_fixed(np.float32(0.125), 2) == "0.13"      # half up
    assert E.zig_fixed(np.float32(12.3456789), 4) == "12.3457" and E.zig_fixed(np.float32(99.995), 2) == "100.00"
    assert E.zig_fixed(np.float32(2.0), 0) == "2" and E.zig_fixed(np.float32(-0.001), 1) == "-0.0"


def _random_sets(rng, n_sets, n_refs):
    refs, sets = [], []
    for _ in range(n_refs):
        t, lab = 0.0, []
        for _ in range(rng.integers(0, 12)):
            t += rng.uniform(0.2, 30)
            d = rng.uniform(0.2, 9)
            lab.append((np.float32(t), np.float32(t + d)))
            t += d
        refs.append(np.array(lab, dtype=[("from_sec", "<f4"), ("to_sec", "<f4")]))
    for _ in range(n_sets):
        t, segs = 0, []
        for _ in range(rng.integers(0, 14)):
            t += int(rng.uniform(0.1, 25) * 48000)
            d = int(rng.uniform(0.7, 12) * 48000)
            segs.append((t, t + d, rng.uniform(0, 1), rng.uniform(0, 1)))
            t += d - int(rng.uniform(0, 0.5) * 48000) * int(rng.integers(0, 2))      # sometimes overlapping margins
        sets.append(segs)
    return refs, sets


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [_cfg(), _cfg(ignore_shorter_than_sec=0.7, extrude_start=5, extrude_end=10, fill_gaps=5),
                                 _cfg(ignore_shorter_than_sec=2.0, extrude_start=0.5, extrude_end=0.25, fill_gaps=40)])
def test_eval_stats_batch_bit_exact(fv, cfg):
    from fvad import evaluator as ev
    rng = np.random.default_rng(7)
    n_sets, n_refs = 300, 9
    refs, sets = _random_sets(rng, n_sets, n_refs)
    per = [np.array(s, dtype=fv.SEG_DTYPE) if s else np.zeros(0, fv.SEG_DTYPE) for s in sets]
    segs, counts = ev.stack_segments(per)
    set_ref = rng.integers(0, n_refs, n_sets).astype(np.uint32)
    got = ev.stats_batch(segs, counts, set_ref, refs, 48000, ev.StatConfig(*[cfg[k] for k in ("ignore_shorter_than_sec", "extrude_start", "extrude_end", "fill_gaps")]))
    nontrivial = 0
    for i in range(n_sets):
        ins = E.vad_segments_to_speech(per[i], 48000)
        rf = [E.Seg(r["from_sec"], r["to_sec"], "ref") for r in refs[set_ref[i]]]
        ins, rf = E.init_and_run(ins, rf)
        want = E.from_evaluator(ins, rf, cfg)
        for k in ev.STATS_DTYPE.names:
            g, w = got[i][k], np.float32(want[k])
            assert (np.isnan(g) and np.isnan(w)) or g.tobytes() == w.tobytes(), (i, k, g, w)
        nontrivial += int(want["true_positives_sec"] > 0 and want["false_positives_sec"] > 0 and want["false_negatives_sec"] > 0)
    assert nontrivial > 30


def test_host_formats_against_oracle(fv):
    """Audacity text and the label parser of the product host code equal the oracle restatement's."""
    from fvad import evaluator as ev
    rng = np.random.default_rng(3)
    refs, sets = _random_sets(rng, 6, 3)
    for i, s in enumerate(sets):
        seg = np.array(s, dtype=fv.SEG_DTYPE) if s else np.zeros(0, fv.SEG_DTYPE)
        ref = refs[i % 3]
        ins = E.vad_segments_to_speech(seg, 48000)
        rf = [E.Seg(r["from_sec"], r["to_sec"], "ref") for r in ref]
        ins, rf = E.init_and_run(ins, rf)
        want = E.serialize_audacity(ins, rf)
        got = ev.serialize_audacity(ev.vad_to_seconds(seg, 48000), ev.debug_info(seg), ref)
        assert got == want
        back = ev.parse_audacity_segments(got)
        oback = E.parse_audacity(want)
        assert len(back) == len(oback) and all(b["from_sec"] == o.from_sec and b["to_sec"] == o.to_sec for b, o in zip(back, oback))
