"""CPU restatement of the reference's Evaluator and simulator bookkeeping.  TEST INFRASTRUCTURE ONLY: imported by
tests/ (never by the product under formula-vad_b200/).  Plain Python over numpy float32 scalars, one statement per
reference statement; pinned by the reference's own unit tests `calcFalsePositiveSec #1/#2`
(src/Evaluator/statistics.zig:286-360), which tests/test_evaluator.py replays.

Follows: src/Evaluator.zig:90-156 (initAndRun), src/Evaluator/SpeechSegment.zig:18-56,
src/Evaluator/statistics.zig:85-284, src/Evaluator/formats.zig:7-56, src/simulator/SimulationInstance.zig:218-243
(storeResult), src/simulator/report_generator.zig:10-116."""
from decimal import Decimal, ROUND_HALF_UP

import numpy as np

f32 = np.float32


class Seg:                                              # SpeechSegment.zig:7-16
    def __init__(self, from_sec, to_sec, side, debug_info=None):
        self.id = -1
        self.from_sec, self.to_sec, self.side = f32(from_sec), f32(to_sec), side
        self.opposite = None
        self.debug_info = debug_info

    def duration(self):                                 # :18-20
        return f32(self.to_sec - self.from_sec)

    def overlap_with(self, o):                          # :22-27
        return f32(min(self.to_sec, o.to_sec) - max(self.from_sec, o.from_sec))

    def has_match(self):                                # :29-31
        return len(self.opposite) > 0

    def to_comment(self):                               # :33-39
        return (self.debug_info or "") if self.has_match() else "UNMATCHED " + (self.debug_info or "")


def find_overlapping(target, others):                   # SpeechSegment.zig:41-51
    return [o for o in others if target.overlap_with(o) > f32(0.0)]


def init_and_run(inputs, refs):                         # Evaluator.zig:90-156 (inputs / refs already ordered by start)
    inputs, refs = list(inputs), list(refs)
    for i, s in enumerate(inputs):
        s.id = i + 1
    for i, s in enumerate(refs):
        s.id = i + 1
    for s in inputs:
        s.opposite = find_overlapping(s, refs)
    for r in refs:
        r.opposite = find_overlapping(r, inputs)
    return inputs, refs


def extrude_segments(segments, cfg):                    # statistics.zig:225-250
    cl = [Seg(s.from_sec, s.to_sec, s.side) for s in segments]
    if not cl:
        return cl
    cl[0].from_sec = f32(cl[0].from_sec - f32(cfg["extrude_start"]))
    cl[-1].to_sec = f32(cl[-1].to_sec + f32(cfg["extrude_end"]))
    for i in range(len(cl) - 1):
        if f32(cl[i + 1].from_sec - cl[i].to_sec) <= f32(cfg["fill_gaps"]):
            cl[i].to_sec = cl[i + 1].from_sec
    return cl


def calc_overlap_many(seg, others):                     # statistics.zig:279-283
    ov = f32(0.0)
    for o in others:
        ov = f32(ov + max(f32(0.0), seg.overlap_with(o)))
    return ov


def calc_false_positive_sec(vad_seg, cfg):              # statistics.zig:193-205
    assert vad_seg.side == "vad"
    return f32(vad_seg.duration() - calc_overlap_many(vad_seg, extrude_segments(vad_seg.opposite, cfg)))


def calc_true_positive_sec(vad_seg, cfg):               # statistics.zig:207-216
    return f32(vad_seg.duration() - calc_false_positive_sec(vad_seg, cfg))


def calc_false_negative_sec(ref_seg, cfg):              # statistics.zig:218-229
    assert ref_seg.side == "ref"
    return f32(ref_seg.duration() - calc_overlap_many(ref_seg, ref_seg.opposite))


def f_score(beta, precision, recall):                   # statistics.zig:171-173
    beta = f32(beta)
    with np.errstate(all="ignore"):
        return f32(f32(f32(f32(1) + f32(beta * beta)) * f32(precision * recall)) / f32(f32(f32(beta * beta) * precision) + recall))


def fm_index(precision, recall):                        # statistics.zig:176-178
    with np.errstate(all="ignore"):
        return f32(np.sqrt(f32(precision * recall)))


def from_evaluator(inputs, refs, cfg):                  # statistics.zig:85-113
    st = dict(total_positives_sec=f32(0), true_positives_sec=f32(0), false_positives_sec=f32(0), false_negatives_sec=f32(0))
    for s in inputs:
        st["false_positives_sec"] = f32(st["false_positives_sec"] + calc_false_positive_sec(s, cfg))
        tp = calc_true_positive_sec(s, cfg)
        st["true_positives_sec"] = f32(st["true_positives_sec"] + tp)
        st["total_positives_sec"] = f32(st["total_positives_sec"] + tp)
    for r in refs:
        if r.duration() < f32(cfg["ignore_shorter_than_sec"]):
            continue
        fn = calc_false_negative_sec(r, cfg)
        st["false_negatives_sec"] = f32(st["false_negatives_sec"] + fn)
        st["total_positives_sec"] = f32(st["total_positives_sec"] + fn)
    with np.errstate(all="ignore"):
        st["true_positive_rate"] = f32(st["true_positives_sec"] / st["total_positives_sec"])
        st["false_negative_rate"] = f32(st["false_negatives_sec"] / st["total_positives_sec"])
        st["false_discovery_rate"] = f32(st["false_positives_sec"] / f32(st["false_positives_sec"] + st["true_positives_sec"]))
        st["precision"] = f32(st["true_positives_sec"] / f32(st["true_positives_sec"] + st["false_positives_sec"]))
    st["f_score_beta"] = f32(0.7)
    st["f_score"] = f_score(st["f_score_beta"], st["precision"], st["true_positive_rate"])
    st["fm_index"] = fm_index(st["precision"], st["true_positive_rate"])
    return st


def vad_segments_to_speech(vad_segments, sample_rate):  # SimulationInstance.storeResult, :218-243
    out = []
    for s in vad_segments:
        frm = f32(f32(int(s["sample_from"])) / f32(sample_rate))
        to = f32(f32(int(s["sample_to"])) / f32(sample_rate))
        dbg = "rnn:%s%% vr:%s" % (zig_fixed(f32(f32(s["rnn_vad"]) * f32(100)), 2), zig_fixed(f32(s["vol_ratio"]), 2))
        out.append(Seg(frm, to, "vad", dbg))
    return out


def zig_fixed(x, prec):
    """Zig 0.11 `{d:.N}` of an f32: the shortest round-trip decimal digits of the f32, then round-half-up at N places
    (std.fmt.formatFloatDecimal over errol3 output)."""
    x = f32(x)
    if np.isnan(x):
        return "nan"
    if np.isinf(x):
        return "inf" if x > 0 else "-inf"
    d = Decimal(np.format_float_positional(x, unique=True, trim="-"))
    q = d.quantize(Decimal(1).scaleb(-prec), rounding=ROUND_HALF_UP)
    s = format(q, "f")
    if q == 0 and d < 0 and not s.startswith("-"):
        s = "-" + s
    return s


def parse_audacity(txt):                                # formats.zig:7-36 (note: it splits the ORIGINAL text, "\r" kept)
    segs = []
    for line in txt.split("\n"):
        fields = line.split("\t")
        if len(fields) < 2:
            continue
        segs.append(Seg(f32(float(fields[0])), f32(float(fields[1])), "ref"))
    return segs


def serialize_audacity(inputs, refs):                   # formats.zig:38-56
    out = []
    for s in inputs:
        out.append("%s\t%s\t%s\n" % (zig_fixed(s.from_sec, 4), zig_fixed(s.to_sec, 4), s.to_comment()))
    for r in refs:
        if r.has_match():
            continue
        out.append("%s\t%s\t%s\n" % (zig_fixed(r.from_sec, 4), zig_fixed(r.to_sec, 4), "missed"))
    return "".join(out)
