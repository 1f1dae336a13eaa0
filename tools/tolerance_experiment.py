"""VERDICT round 1, Next #10: what would a tensor-core (reassociated / reduced-precision) RNN do to the RESULTS?

Runs the pipeline with knob rnn_tolerance_mode = 0 (the exact product path), 1 (f32, four partial sums), 2 (TF32
activations), 3 (BF16 activations) -- rn_core.cuh rnn_chain_tol emulates the arithmetic, not the speed -- and reports
  (a) on S streams x SHORT s with the denoised PCM and the per-frame VAD probability traced: max / mean absolute
      difference from mode 0 (PCM in normalised [-1, 1] units and in the RNNoise int16 scale),
  (b) on configs[1] (S streams x LONG s): how many streams keep a segment list IDENTICAL to mode 0, and how many
      segments exist / moved.
usage: tolerance_experiment.py [S=20] [SHORT=300] [LONG=7200]      -> one JSON object on stdout"""
import sys, os, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
import numpy as np, torch, fvad

S = int(sys.argv[1]) if len(sys.argv) > 1 else 20
SHORT = int(sys.argv[2]) if len(sys.argv) > 2 else 300
LONG = int(sys.argv[3]) if len(sys.argv) > 3 else 7200
SEED = 0xF1AD5EED00000000
MODES = {0: "exact (in-order f32)", 1: "f32, four partial sums", 2: "tf32 activations", 3: "bf16 activations"}


def run(mode, secs, trace):
    fvad.debug_set("rnn_tolerance_mode", str(mode))
    n = secs * 48000
    x = torch.empty(S, n, dtype=torch.float32, device="cuda")
    fvad.synth_device(x.data_ptr(), n, SEED, 0, S, 0, n)
    torch.cuda.synchronize()
    b = fvad.PipelineBatch(S, trace=trace)
    t0 = time.time()
    b.push_device(x.data_ptr(), n, n, n)
    b.sync()
    dt = time.time() - t0
    out = {"segs": [b.segments(s).copy() for s in range(S)], "wall_s": dt}
    if trace:
        out["pcm"] = [b.trace_denoised(s) for s in range(S)]
        out["vad"] = [b.trace_rnn_vad(s) for s in range(S)]
    b.close()
    del x
    torch.cuda.empty_cache()
    return out


def seg_key(a):
    return [(int(r["sample_from"]), int(r["sample_to"])) for r in a]


res = {"streams": S, "short_s": SHORT, "long_s": LONG, "modes": {}}
short0 = run(0, SHORT, fvad.TRACE_DENOISED | fvad.TRACE_RNN_VAD)
long0 = run(0, LONG, 0)
res["modes"]["0"] = {"what": MODES[0], "segments_long": sum(len(a) for a in long0["segs"]), "wall_long_s": long0["wall_s"]}
for mode in (1, 2, 3):
    sh = run(mode, SHORT, fvad.TRACE_DENOISED | fvad.TRACE_RNN_VAD)
    pcm_max = max(float(np.max(np.abs(a - b))) for a, b in zip(sh["pcm"], short0["pcm"]))
    pcm_mean = float(np.mean([np.mean(np.abs(a - b)) for a, b in zip(sh["pcm"], short0["pcm"])]))
    ref_max = max(float(np.max(np.abs(b))) for b in short0["pcm"])
    vad_max = max(float(np.max(np.abs(a - b))) for a, b in zip(sh["vad"], short0["vad"]))
    vad_mean = float(np.mean([np.mean(np.abs(a - b)) for a, b in zip(sh["vad"], short0["vad"])]))
    frac_pcm_equal = float(np.mean([np.mean(a == b) for a, b in zip(sh["pcm"], short0["pcm"])]))
    del sh
    lg = run(mode, LONG, 0)
    same = 0; moved = 0; count_diff = 0
    for a, b in zip(lg["segs"], long0["segs"]):
        ka, kb = seg_key(a), seg_key(b)
        if ka == kb:
            same += 1
        else:
            count_diff += abs(len(ka) - len(kb))
            moved += len(set(ka) ^ set(kb))
    res["modes"][str(mode)] = {
        "what": MODES[mode],
        "pcm_abs_diff_max": pcm_max, "pcm_abs_diff_max_int16_scale": pcm_max * 32767.0, "pcm_abs_diff_mean": pcm_mean,
        "pcm_abs_max_of_reference": ref_max, "pcm_samples_bit_equal_frac": frac_pcm_equal,
        "vad_prob_abs_diff_max": vad_max, "vad_prob_abs_diff_mean": vad_mean,
        "within_1e-4": bool(pcm_max <= 1e-4 and vad_max <= 1e-4),
        "streams_with_identical_segment_list": same, "streams": S,
        "segments_long": sum(len(a) for a in lg["segs"]), "segment_boundaries_differing": moved,
        "segment_count_difference": count_diff, "wall_long_s": lg["wall_s"],
    }
    print(json.dumps({mode: res["modes"][str(mode)]}), file=sys.stderr, flush=True)
fvad.debug_set("rnn_tolerance_mode", "0")
print(json.dumps(res))
