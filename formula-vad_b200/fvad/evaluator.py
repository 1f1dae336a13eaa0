"""Evaluator drop-in (src/Evaluator.zig, src/Evaluator/{SpeechSegment,statistics,formats}.zig,
src/simulator/report_generator.zig) on top of the C ABI: the per-result-set statistics (TP/FP/FN seconds with
extrusion and gap filling) run on the GPU for any number of (stream, config) result sets at once
(`fvad_eval_stats_batch`, one thread per set); parsing / serialising Audacity label files, the aggregate and the
report text are host code, as in the reference.  SURVEY.md 8(f) N1/N3."""
import ctypes as C
from decimal import Decimal, ROUND_HALF_UP

import numpy as np

from . import SEG_DTYPE, FvadError, _check, lib  # noqa: F401

STATS_DTYPE = np.dtype([(k, "<f4") for k in (
    "total_positives_sec", "true_positives_sec", "false_positives_sec", "false_negatives_sec", "true_positive_rate",
    "false_negative_rate", "false_discovery_rate", "precision", "fm_index", "f_score", "f_score_beta")])
REF_DTYPE = np.dtype([("from_sec", "<f4"), ("to_sec", "<f4")])


class StatConfig(C.Structure):
    """statistics.StatConfig (statistics.zig:77-83)."""
    _fields_ = [("ignore_shorter_than_sec", C.c_float), ("extrude_start", C.c_float), ("extrude_end", C.c_float),
                ("fill_gaps", C.c_float)]


def simulator_stat_config(min_vad_duration_sec):
    """The StatConfig simulator.main uses (src/simulator.zig:121-126)."""
    return StatConfig(min_vad_duration_sec, 5, 10, 5)


# ---- number formatting ------------------------------------------------------------------------------------
def zig_fmt(x, prec, width=0):
    """Zig 0.11 `{d: >W.P}` of an f32: shortest round-trip digits, rounded half-up at P places, right-aligned to W."""
    x = np.float32(x)
    if np.isnan(x):
        s = "nan"
    elif np.isinf(x):
        s = "inf" if x > 0 else "-inf"
    else:
        d = Decimal(np.format_float_positional(x, unique=True, trim="-"))
        q = d.quantize(Decimal(1).scaleb(-prec), rounding=ROUND_HALF_UP)
        s = format(q, "f")
        if q == 0 and d < 0 and not s.startswith("-"):
            s = "-" + s
    return s.rjust(width)


# ---- Audacity label files (formats.zig) -------------------------------------------------------------------
def parse_audacity_segments(txt):
    """formats.parseAudacitySegments (formats.zig:7-36): `from<TAB>to[<TAB>comment]` per line -> REF_DTYPE array."""
    rows = []
    for line in txt.split("\n"):
        fields = line.split("\t")
        if len(fields) < 2:
            continue
        rows.append((np.float32(float(fields[0])), np.float32(float(fields[1]))))
    return np.array(rows, dtype=REF_DTYPE)


def read_parse_audacity_segments(path):
    with open(path, "r", newline="") as f:              # Evaluator.readParseAudacitySegments, Evaluator.zig:176-183
        return parse_audacity_segments(f.read())


def sort_by_start(a):
    """Evaluator.initAndRun orders both sides by start (Evaluator.zig:108-109)."""
    return a[np.argsort(a["from_sec"], kind="stable")]


def vad_to_seconds(segments, sample_rate):
    """SimulationInstance.storeResult (SimulationInstance.zig:226-227): f32(sample) / f32(sample_rate)."""
    out = np.empty(len(segments), dtype=REF_DTYPE)
    sr = np.float32(sample_rate)
    out["from_sec"] = segments["sample_from"].astype(np.float32) / sr
    out["to_sec"] = segments["sample_to"].astype(np.float32) / sr
    return out


def debug_info(segments):
    """`rnn:{d:.2}% vr:{d:.2}` (SimulationInstance.zig:229-233)."""
    return ["rnn:%s%% vr:%s" % (zig_fmt(np.float32(s["rnn_vad"]) * np.float32(100), 2), zig_fmt(s["vol_ratio"], 2)) for s in segments]


def _overlap_matrix(a, b):
    lo = np.maximum(a["from_sec"][:, None], b["from_sec"][None, :])
    hi = np.minimum(a["to_sec"][:, None], b["to_sec"][None, :])
    return (hi - lo).astype(np.float32)                 # SpeechSegment.overlapWith, SpeechSegment.zig:22-27


def serialize_audacity(vad_sec, comments, ref_sec):
    """formats.serializeEvaluatorToAudacityTxt (formats.zig:38-56): matched / UNMATCHED inputs, then missed labels."""
    ov = _overlap_matrix(vad_sec, ref_sec) > 0 if len(vad_sec) and len(ref_sec) else np.zeros((len(vad_sec), len(ref_sec)), bool)
    out = []
    for i, s in enumerate(vad_sec):
        c = comments[i] if comments is not None else ""
        if not ov[i].any():
            c = "UNMATCHED " + c                        # SpeechSegment.toComment, SpeechSegment.zig:33-39
        out.append("%s\t%s\t%s\n" % (zig_fmt(s["from_sec"], 4), zig_fmt(s["to_sec"], 4), c))
    for j, r in enumerate(ref_sec):
        if ov[:, j].any():
            continue
        out.append("%s\t%s\t%s\n" % (zig_fmt(r["from_sec"], 4), zig_fmt(r["to_sec"], 4), "missed"))
    return "".join(out)


# ---- statistics on the device -------------------------------------------------------------------------------
def stats_batch(segs, counts, set_ref, refs, sample_rate, config, device=-1):
    """statistics.fromEvaluator for every result set at once.
    segs: SEG_DTYPE [n_sets, max_segs] (fvad_vadm_sweep's layout, or stacked per-stream lists); counts [n_sets];
    set_ref [n_sets] -> index into refs, a list of REF_DTYPE arrays ordered by start.  Returns STATS_DTYPE [n_sets]."""
    segs = np.ascontiguousarray(segs, dtype=SEG_DTYPE)
    n_sets, max_segs = segs.shape
    counts = np.ascontiguousarray(counts, np.uint32)
    set_ref = np.ascontiguousarray(set_ref, np.uint32)
    offs = np.zeros(len(refs) + 1, np.uint32)
    offs[1:] = np.cumsum([len(r) for r in refs])
    allr = np.concatenate([np.asarray(r, dtype=REF_DTYPE) for r in refs]) if len(refs) else np.zeros(0, REF_DTYPE)
    rf = np.ascontiguousarray(allr["from_sec"]) if len(allr) else np.zeros(1, np.float32)
    rt = np.ascontiguousarray(allr["to_sec"]) if len(allr) else np.zeros(1, np.float32)
    out = np.empty(n_sets, STATS_DTYPE)
    _check(lib().fvad_eval_stats_batch(segs.ctypes.data, n_sets, max_segs, counts.ctypes.data, set_ref.ctypes.data, rf.ctypes.data,
                                       rt.ctypes.data, offs.ctypes.data, len(refs), sample_rate, C.addressof(config), device,
                                       out.ctypes.data))
    return out


def stats_batch_device(d_segs, n_sets, max_segs, d_counts, set_ref, refs, sample_rate, config):
    """The same with the result sets already on the device (PipelineBatch.sweep's device form, fvad_vadm_sweep_static):
    d_segs / d_counts are device addresses; runs on the current device."""
    set_ref = np.ascontiguousarray(set_ref, np.uint32)
    offs = np.zeros(len(refs) + 1, np.uint32)
    offs[1:] = np.cumsum([len(r) for r in refs])
    allr = np.concatenate([np.asarray(r, dtype=REF_DTYPE) for r in refs]) if len(refs) else np.zeros(0, REF_DTYPE)
    rf = np.ascontiguousarray(allr["from_sec"]) if len(allr) else np.zeros(1, np.float32)
    rt = np.ascontiguousarray(allr["to_sec"]) if len(allr) else np.zeros(1, np.float32)
    out = np.empty(n_sets, STATS_DTYPE)
    _check(lib().fvad_eval_stats_batch_device(C.c_void_p(d_segs), n_sets, max_segs, C.c_void_p(d_counts), set_ref.ctypes.data,
                                              rf.ctypes.data, rt.ctypes.data, offs.ctypes.data, len(refs), sample_rate,
                                              C.addressof(config), out.ctypes.data))
    return out


def stack_segments(per_set):
    """Lists of SEG_DTYPE arrays -> ([n_sets, max_segs] array, counts)."""
    m = max([len(s) for s in per_set] + [1])
    segs = np.zeros((len(per_set), m), SEG_DTYPE)
    for i, s in enumerate(per_set):
        segs[i, :len(s)] = s
    return segs, np.array([len(s) for s in per_set], np.uint32)


def aggregate(stats):
    """statistics.aggregate (statistics.zig:115-168): f32 sums in instance order."""
    f = np.float32
    agg = {k: f(0) for k in ("total_positives_sec", "true_positives_sec", "false_positives_sec", "false_negatives_sec")}
    rates = {k: {"min": f(2), "max": f(-2), "sum": f(0)} for k in ("true_positive_rate", "false_negative_rate", "false_discovery_rate", "precision")}
    for s in stats:
        for k in agg:
            agg[k] = f(agg[k] + s[k])
        for k, r in rates.items():
            r["sum"] = f(r["sum"] + s[k])
            if s[k] < r["min"]:
                r["min"] = f(s[k])
            if s[k] > r["max"]:
                r["max"] = f(s[k])
    n = f(len(stats))
    with np.errstate(all="ignore"):
        rates["true_positive_rate"]["overall"] = f(agg["true_positives_sec"] / agg["total_positives_sec"])
        rates["false_negative_rate"]["overall"] = f(agg["false_negatives_sec"] / agg["total_positives_sec"])
        rates["false_discovery_rate"]["overall"] = f(agg["false_positives_sec"] / f(agg["false_positives_sec"] + agg["true_positives_sec"]))
        rates["precision"]["overall"] = f(agg["true_positives_sec"] / f(agg["true_positives_sec"] + agg["false_positives_sec"]))
        for r in rates.values():
            r["avg"] = f(r["sum"] / n)
        beta = f(0.7)
        p, t = rates["precision"]["overall"], rates["true_positive_rate"]["overall"]
        agg["f_score_beta"] = beta
        agg["f_score"] = f(f(f(f(1) + f(beta * beta)) * f(p * t)) / f(f(f(beta * beta) * p) + t))
        agg["fm_index"] = f(np.sqrt(f(p * t)))
    agg.update(rates)
    return agg


DEFINITIONS = (
    "P   (Positives):                            Total duration of real speech segments (from reference labels)\n"
    "TP  (True positives):                       Duration of correctly detected speech segments\n"
    "FP  (False positives):                      Duration of incorrectly detected speech segments\n"
    "FN  (False negatives):                      Duration of missed speech segments\n"
    "TPR (True positive rate, sensitivity):      Probability that VAD detects a real speech segment. = TP / P \n"
    "FNR (False negative rate, miss rate):       Probability that VAD misses a speech segment.       = FN / P \n"
    "PPV (Precision, Positive predictive value): Probability that detected speech segment is true.   = TP / (TP + FP) \n"
    "FDR (False discovery rate):                 Probability that detected speech segment is false.  = FP / (TP + FP) ")


def simulation_report(names, stats):
    """report_generator.bufPrintSimulationReport (report_generator.zig:29-116)."""
    f = np.float32
    hdr = ("Name", "P", "TP", "FP", "FN", "TPR", "FNR", "PPV", "FDR (!)")
    w = (30, 4, 4, 4, 4, 6, 6, 6, 8)
    out = ["\n\n=> Definitions\n\n%s" % DEFINITIONS, "\n\n=> Performance Report\n\n"]
    out.append("| " + " | ".join(h.rjust(n) for h, n in zip(hdr, w)) + " |\n")
    out.append("| " + " | ".join("-" * n for n in w) + " |\n")
    for name, s in zip(names, stats):
        cells = [name.rjust(30)] + [zig_fmt(s[k], 0, 4) for k in ("total_positives_sec", "true_positives_sec", "false_positives_sec", "false_negatives_sec")]
        cells += [zig_fmt(f(s[k]) * f(100), 1, 5) + "%" for k in ("true_positive_rate", "false_negative_rate", "precision")]
        cells += [zig_fmt(f(s["false_discovery_rate"]) * f(100), 1, 7) + "%"]
        out.append("| " + " | ".join(cells) + " |\n")
    a = aggregate(stats)
    out.append("\n=> Aggregate stats \n\n")
    out.append("Total speech duration  (P): %s sec\n" % zig_fmt(a["total_positives_sec"], 1, 7))
    out.append("True positives        (TP): %s sec\n" % zig_fmt(a["true_positives_sec"], 1, 7))
    out.append("False positives       (FP): %s sec\n" % zig_fmt(a["false_positives_sec"], 1, 7))
    out.append("False negatives       (FN): %s sec" % zig_fmt(a["false_negatives_sec"], 1, 7))
    out.append("    Min.    Avg.    Max. \n")
    for label, k in (("True positive rate   (TPR)", "true_positive_rate"), ("False negative rate  (FNR)", "false_negative_rate"),
                     ("Precision            (PPV)", "precision"), ("False discovery rate (FDR)", "false_discovery_rate")):
        r = a[k]
        out.append("%s:   %s%%  |  %s%% /%s%% /%s%% \n" % (label, zig_fmt(r["overall"] * f(100), 1, 5), zig_fmt(r["min"] * f(100), 1, 5),
                                                            zig_fmt(r["avg"] * f(100), 1, 5), zig_fmt(r["max"] * f(100), 1, 5)))
    out.append("F-Score (β = %s)       :   %s%% \n" % (zig_fmt(a["f_score_beta"], 2, 5), zig_fmt(a["f_score"] * f(100), 1, 5)))
    out.append("Fowlkes-Mallows index     :   %s%% \n" % zig_fmt(a["fm_index"] * f(100), 1, 5))
    return "".join(out)
