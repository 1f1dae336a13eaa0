#!/usr/bin/env python3
"""bench.py -- the driver-facing benchmark of the Formula-VAD AudioPipeline hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Metric (BASELINE.json): audio-seconds processed per wall second (x real-time).  One STEP = one pass of
the whole hot path (RNNoise denoise + VAD probability -> 480->2048 re-framing -> windowed STFT + speech-band
sum -> VADMachine -> segment list) over one batch of synthetic input:

  N = 1   BASELINE.json configs[1]: 20 synthetic 48 kHz mono driver streams x 2 h, default plan.
  N > 1   the same 20 streams x 2 h PER GPU (weak scaling; N = 8 is configs[3]: 160 streams x 2 h sharded
          by stream), one process per GPU under torchrun, no data-path collective, one final gather of the
          segment lists to rank 0 (fvad/shard.py over NCCL) inside every step.

Keys of the one JSON line (rank 0):
`value`   : inputs resident in HBM before the timed region (27.6 GB per GPU >> L2, so no L2 flush is needed).
`e2e`     : the same step through the reference-facing call with HOST buffers (fvad_batch_push = AudioPipeline.
            pushSamples): H2D copies from pinned memory and the D2H segment read are inside the timed region.
`e2e_sim` : the same, pushed 48 000 samples at a time -- the call pattern of the reference's only caller
            (src/simulator/SimulationInstance.zig:194-203) and of the reference arm.
`configs0`: BASELINE.json configs[0]: one stream x 60 s through fvad_pipeline_push, single-thread CPU beside it.
`roofline`: the kernel with the largest share of the step, timed live with CUDA events on its own stream (the library's
            per-launch event profiler).  It is a dependent chain: reported against its latency bound (cycles per sample
            vs the chain's dependent operations), with its HBM figure beside it.
`stft`    : BASELINE.json configs[2] shape (4096 streams, STFT + band sum only), the bandwidth-bound stage, against the
            measured HBM peak of MEASURED_PEAKS.json.
`batched` : north_star's many-stream regime (1184 streams per GPU, every rank): device-resident, and host-fed through
            fvad_batch_push from pinned and (N = 1) pageable rows.
`configs3_strong`: BASELINE.json configs[3] as written: 160 streams x 2 h over the N GPUs (strong scaling).
`sweep`   : BASELINE.json configs[4]: 256 VADMachine configs x 20 streams over the run's own STFT records, on device.
`segments_sha_first20`: hash of the packed segment records of global streams 0..19 -- equal at every N.
`cpu_baseline` / `--impl reference`: the reference's own C sources (oracle/_ref) under the restated Zig glue
            on the box's host cores, one OS thread per stream as src/simulator.zig:217-228 does.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "formula-vad_b200"), ROOT]

SEED = 0xF1AD5EED00000000
METRIC = "audio-sec processed/sec (x real-time)"
UNIT = "audio_s/s"
SR = 48000
_REAL_STDOUT = sys.stdout

# Algorithmic bytes per unit for each kernel (DESIGN.md "Kernels"): the stage's compulsory input + output,
# tables and weights resident.  Unit = one 480-sample RNNoise frame unless noted.
ALGO_BYTES = {
    "k_stft2048<false>": (8196, "stft_frame"),        # 2048 f32 in + one band sum out
    "k_stft2048<true>": (12292, "stft_frame"),        # + 1025 magnitudes out
    "k_rn_biquad": (3840, "frame"),                   # 480 f32 in, 480 f32 out
    "k_rn_front": (1920 + 4 * (42 + 2), "frame"),
    "k_rn_analysis": (2008, "frame"),                 # 480 f32 in + 22 band energies out
    "k_rn_pitch": (1920 + 200, "frame"),
    "k_rn_pfeat": (1920 + 3848 + 88, "frame"),
    "k_rn_rnn": (4 * (42 + 23), "frame"),             # 42 features in, 22 gains + vad out; weights in smem
    "k_rn_synth": (2 * 3848 + 1920, "frame"),
    "k_rn_out": (3 * 1920, "frame"),
    "k_vadm": (12, "stft_frame"),
}
BIQUAD_CHAIN_OPS = 7        # dependent f32 operations per sample on the chain lane (rn_core.cuh biquad_chain_step)
F32_OP_CYCLES = 4           # dependent-issue latency of FADD / FMUL / FFMA on sm_100 (B300_MICROARCH.md)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=20, help="streams per GPU (default: BASELINE configs[1])")
    ap.add_argument("--seconds", type=int, default=7200, help="audio seconds per stream (default 2 h)")
    ap.add_argument("--ref-sample-seconds", type=int, default=240, help="--impl reference: seconds per stream per step")
    ap.add_argument("--cpu-sample-seconds", type=int, default=60, help="cpu_baseline: seconds per stream")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sim", action="store_true", help="skip the 48000-sample-push leg")
    ap.add_argument("--no-stft", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the many-stream leg")
    ap.add_argument("--no-strong", action="store_true", help="skip configs[3] as written (160 streams over N GPUs)")
    ap.add_argument("--no-sweep", action="store_true", help="skip configs[4] (256-config sweep)")
    ap.add_argument("--batched-streams", type=str, default="1184", help="streams per GPU of the many-stream leg; a comma list "
                    "gives a device-resident sweep table (the first entry is the one fed from the host)")
    ap.add_argument("--batched-seconds", type=int, default=150)
    ap.add_argument("--strong-streams", type=int, default=160)
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def mem_available():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 0


def sha_first(records, n_streams=20):
    """Hash of the packed records (fvad.shard.WIRE_DTYPE) of global streams 0..n_streams-1, in (stream, machine, time) order."""
    if records is None:
        return None
    sel = records[records["stream"] < n_streams]
    return hashlib.sha256(sel.tobytes()).hexdigest()[:16]


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's C libraries + restated Zig glue, one OS thread per stream.  Imports nothing of the product:
# the input comes from oracle/libsynth_host.so (the same generator, bit-equal: tests/test_oracle.py).
# ------------------------------------------------------------------------------------------------
def cpu_kind():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from oracle import have_reference, have_reference_fast
    if have_reference_fast():
        return "reference_fast", "reference", "gcc -O3 -march=x86-64-v3, FMA contraction allowed (stand-in for zig ReleaseFast, unavailable on this box)"
    if have_reference():
        return "reference", "reference", "gcc -O2 -ffp-contract=off (the parity build; zig ReleaseFast unavailable on this box)"
    return "port", "port", "gcc -O2 -ffp-contract=off, plain-C restatement (oracle/_ref not prebuilt)"


def cpu_run(n_streams, seconds, faithful=True, first_stream=0, threads=None):
    """One pass: n_streams pipelines, one OS thread each (src/simulator.zig:217-228 spawns a thread per instance and lets
    the OS time-slice them), 48000-sample pushes (SimulationInstance.zig:194-203).  Returns a dict."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from concurrent.futures import ThreadPoolExecutor
    import oracle
    load, kind, compiler = cpu_kind()
    orc = oracle.Oracle(load)
    threads = threads or n_streams
    n = seconds * SR
    with ThreadPoolExecutor(min(threads, os.cpu_count() or 1)) as ex:           # input generation is outside the timed region
        pcm = list(ex.map(lambda s: oracle.synth(SEED, first_stream + s, 0, n), range(n_streams)))
    with ThreadPoolExecutor(threads) as ex:
        def one(s):
            t = time.perf_counter()
            r = orc.run_pipeline([pcm[s]], faithful=faithful, push=48000)
            return len(r["segments"][0]), time.perf_counter() - t
        t0 = time.perf_counter()
        res = list(ex.map(one, range(n_streams)))
        wall = time.perf_counter() - t0
    per = sorted(w for _, w in res)
    return {"value": n_streams * seconds / wall, "wall_s": wall, "threads": threads, "kind": kind, "compiler": compiler,
            "segments": int(sum(c for c, _ in res)), "per_stream_wall_s": {"min": per[0], "median": per[len(per) // 2], "max": per[-1]},
            "host_cpus": os.cpu_count()}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # the WHOLE job's streams (weak scaling: streams-per-GPU x N), on the host cores of the box, by rank 0 alone
    S, secs = args.streams * max(1, args.gpus), args.ref_sample_seconds
    for _ in range(args.warmup):
        cpu_run(S, max(10, secs // 8))
    walls, r = [], None
    for _ in range(args.steps):
        r = cpu_run(S, secs)
        walls.append(r["wall_s"])
    wall = sum(walls) / len(walls)
    value = S * secs / wall
    sample = (f"first {secs} s of each of the {S} streams per step (a full step is {args.seconds} s per stream: a rate, so the "
              f"sample scales; full-length equivalent {S * args.seconds / value:.0f} s per step), one OS thread per stream as "
              f"src/simulator.zig:217-228 ({r['threads']} threads on {r['host_cpus']} CPUs), 48000-sample pushes, KissFFT state "
              f"regenerated per FFT as src/FFT.zig:84")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (seeded generator, same streams as the b200 arm; oracle/libsynth_host.so)",
        "config": {"workload": f"configs[1]: {args.streams} synthetic 48 kHz mono streams x 2 h per GPU, RNNoise denoiser + VAD, "
                               f"default plan" + ("" if args.gpus <= 1 else f"; {S} streams, the {args.gpus}-GPU job's whole workload (configs[3] at N=8)"),
                   "streams": S, "streams_per_gpu": args.streams, "seconds_per_stream": args.seconds,
                   "sample_seconds_per_stream": secs, "push": 48000},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(r["threads"], r["host_cpus"] or 1), "threads": r["threads"],
                         "host_cpus": r["host_cpus"], "kind": r["kind"], "sample": sample, "compiler": r["compiler"],
                         "per_stream_wall_s": r["per_stream_wall_s"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "segments": r["segments"],
    }
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# clocks sampling
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def b200_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import fvad
    from fvad import shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    L = fvad.lib()
    if L.fvad_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device -- libfvad_cuda has no CPU fallback: " + L.fvad_last_error().decode())
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        # a leg that one rank skips and another does not would otherwise hang for NCCL's default ten minutes
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=240))
    S, secs = args.streams, args.seconds
    n = secs * SR
    first_stream = rank * S                      # weak scaling: every rank owns its own S streams
    K, W = args.steps, args.warmup
    CH = 600 * SR
    peak, peak_src = peaks()
    errors = {}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def host_mem_per_rank():
        """MemAvailable split over the ranks of this box, agreed on by all of them (every rank must take the same branch)."""
        a = torch.tensor([mem_available()], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(a, op=dist.ReduceOp.MIN)
        return int(a[0]) // world

    def timed(fn, k, ext):
        """k calls of fn between two CUDA events on the batch's own stream, a barrier + synchronize on both sides;
        returns (device ms, wall ms) as the MAX over ranks, and fn's last result."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(ext)
        out = None
        for _ in range(k):
            out = fn()
        e1.record(ext)
        barrier()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = torch.tensor([e0.elapsed_time(e1), wall_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), float(ms[1]), out

    def fill_device(t, first_id, n_streams, n_samples, first_sample=0):
        for off in range(0, n_samples, CH):
            m = min(CH, n_samples - off)
            fvad.synth_device(t.data_ptr() + 4 * off, t.stride(0), SEED, first_id, n_streams, first_sample + off, m)

    def collect_from(batch, n_streams, first_id):
        per = [[batch.segments(s, 0)] for s in range(n_streams)]
        rec = shard.pack_segments(first_id, per)
        allrec = shard.gather_segments(rec, device=dev) if world > 1 else rec
        return rec, allrec

    def host_rows(t, n_rows, off=0):
        base, stride = t.data_ptr(), t.stride(0)
        return [base + 4 * (stride * r + off) for r in range(n_rows)]

    # ---- inputs resident in HBM ---------------------------------------------------------------
    x = torch.empty(S, n, dtype=torch.float32, device=dev)
    fill_device(x, first_stream, S, n)
    torch.cuda.synchronize()
    # the batch keeps its per-STFT-frame features in HBM (12 B per frame per stream) for the configs[4] sweep below
    b = fvad.PipelineBatch(S, device=local, trace=fvad.KEEP_FEATURES)
    ext = torch.cuda.ExternalStream(b.cuda_stream(), device=dev)

    def step_device():
        b.reset()
        b.push_device(x.data_ptr(), n, n, n)
        return collect_from(b, S, first_stream)

    for _ in range(W):
        step_device()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = fvad.launch_count()
    dev_ms, wall_ms, (rec, allrec) = timed(step_device, K, ext)
    launches = fvad.launch_count() - launches0
    # one more step with the library's per-launch CUDA events on (two events around every launch, on the launch's
    # own stream): the per-kernel durations the roofline uses.  Kept out of `value` because the extra event
    # records sit between the kernels of the stage streams.
    fvad.profile_enable(True)
    prof_ms, _, _ = timed(step_device, 1, ext)
    fvad.profile_enable(False)
    prof = fvad.profile_report()
    clocks = sampler.stop()
    ms_per_step = dev_ms / K
    value = world * S * secs / (ms_per_step / 1e3)
    n_segments = int(len(allrec)) if allrec is not None else None
    sha = sha_first(allrec)

    # ---- roofline of the dominant kernel ---------------------------------------------------------
    roofline = None
    shares = {}
    if prof:
        tot = sum(v["total_ms"] for v in prof.values())
        shares = {k: round(v["total_ms"] / tot, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"])}
        name, info = max(prof.items(), key=lambda kv: kv[1]["total_ms"])
        frames = S * (n // 480)                      # the profiled step
        stft_frames = S * (n // 2048)
        per, unit = ALGO_BYTES.get(name, (None, None))
        traffic = None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(name)
        hbm = None
        dur_s = info["total_ms"] / 1e3 / info["launches"]            # average launch duration
        if per is not None:
            units = frames if unit == "frame" else stft_frames
            algo = per * units / info["launches"]                    # bytes per launch
            ach = algo / dur_s / 1e9
            hbm = {"achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                   "algorithmic_bytes_per_launch": algo}
        if name == "k_rn_biquad":
            # a dependent chain per stream: every launch walks (frames per launch per stream) x 480 samples one after the other
            mhz = clocks.get("sm_mhz") or 1965.0
            samples_per_chain = (n // 480) * 480 / info["launches"]
            cyc = dur_s * mhz * 1e6 / samples_per_chain
            bound = BIQUAD_CHAIN_OPS * F32_OP_CYCLES
            roofline = {"kernel": name, "bound": "latency", "achieved": cyc, "peak": bound, "unit": "cycles/sample (lower is better)",
                        "frac": bound / cyc, "cycles_per_sample": cyc, "chain_cycles": bound,
                        "chain": f"{BIQUAD_CHAIN_OPS} dependent f32 operations x {F32_OP_CYCLES} cycles (rn_core.cuh biquad_chain_step)",
                        "sm_mhz": mhz, "avg_launch_ms": 1e3 * dur_s, "launches": info["launches"], "share_of_step": shares[name],
                        "critical_path_share": info["total_ms"] / prof_ms, "traffic": traffic, "hbm": hbm,
                        "note": "sample-sequential IIR per stream: the binding limit is the dependent chain, not HBM or FLOPs "
                                "(DESIGN.md section 5); the bandwidth-bound stage is reported under 'stft'"}
        else:
            roofline = {"kernel": name, "bound": "hbm", "achieved": hbm and hbm["achieved"], "peak": peak, "unit": "GB/s",
                        "frac": hbm and hbm["frac"], "traffic": traffic, "peak_source": peak_src, "avg_launch_ms": 1e3 * dur_s,
                        "launches": info["launches"], "share_of_step": shares[name]}

    # ---- e2e: host buffers through fvad_batch_push --------------------------------------------------
    e2e = e2e_sim = None
    if not args.no_e2e:
        need = S * n * 4
        avail = host_mem_per_rank()
        win = n
        if avail and need * 2 > avail:                            # keep well inside host RAM: bounded window, pushed repeatedly
            win = max(600 * SR, int(avail / 8 / (S * 4)) // (600 * SR) * (600 * SR))
            win = min(win, n)
        h = torch.empty(S, win, dtype=torch.float32, pin_memory=True)
        for off in range(0, win, CH):
            m = min(CH, win - off)
            h[:, off:off + m].copy_(x[:, off:off + m])
        torch.cuda.synchronize()

        def step_host():
            b.reset()
            done = 0
            while done < n:
                m = min(win, n - done)
                b.push_rows(host_rows(h, S), m)
                done += m
            return collect_from(b, S, first_stream)
        step_host()
        e_ms, e_wall, (erec, _) = timed(step_host, K, ext)
        e2e = {"value": world * S * secs / (e_ms / K / 1e3), "unit": UNIT, "h2d_bytes_per_step": S * n * 4,
               "d2h_bytes_per_step": int(erec.nbytes + 8), "ms_per_step": e_ms / K, "wall_ms_per_step": e_wall / K,
               "host_window_s": win // SR, "pinned": True, "push_samples": min(win, n),
               "api": "fvad_batch_push (AudioPipeline.pushSamples drop-in), host PCM in, segment list out"}
        if win == n:
            e2e["segments_equal_device_resident_run"] = len(erec) == len(rec) and bool((erec == rec).all())

        if not args.no_sim:
            # the reference caller's own pattern: one second per call (SimulationInstance.zig:194-203)
            def step_sim():
                b.reset()
                for done in range(0, n, SR):
                    b.push_rows(host_rows(h, S, done % win), min(SR, n - done))
                return collect_from(b, S, first_stream)
            ks = max(1, min(K, 2))
            s_ms, s_wall, (srec, _) = timed(step_sim, ks, ext)
            e2e_sim = {"value": world * S * secs / (s_ms / ks / 1e3), "unit": UNIT, "ms_per_step": s_ms / ks, "wall_ms_per_step": s_wall / ks,
                       "steps": ks, "push_samples": SR, "pushes_per_step": (n + SR - 1) // SR, "pinned": True,
                       "h2d_bytes_per_step": S * n * 4, "d2h_bytes_per_step": int(srec.nbytes + 8),
                       "api": "fvad_batch_push, 48000 samples per call (SimulationInstance.simulateVAD's pattern)"}
            if win == n:
                e2e_sim["segments_equal_device_resident_run"] = len(srec) == len(rec) and bool((srec == rec).all())
            e2e_sim["vs_e2e"] = e2e_sim["value"] / e2e["value"]
        del h

    # ---- configs[4]: the 256-config threshold sweep over this run's own STFT records, on the device ------------------
    sweep = None
    if not args.no_sweep and rank == 0:
        try:
            b.reset()                                              # the features of one full pass of this rank's streams
            b.push_device(x.data_ptr(), n, n, n)                   # (no collective here: the other ranks are not in this leg)
            b.sync()
            sweep = sweep_leg(args, fvad, b, S, secs, first_stream, torch, dev, do_cpu=not args.no_cpu and world == 1)
        except Exception as e:                                     # a leg must not cost the line
            errors["sweep"] = repr(e)
    b.close()
    del x
    torch.cuda.empty_cache()

    # ---- configs[0]: one stream x 60 s through the single-pipeline API ------------------------------------------------
    configs0 = None
    if rank == 0:
        try:
            n0 = 60 * SR
            x0 = fvad.synth_host(SEED, 0, 0, n0)                   # pageable numpy memory, as a caller would hold it
            p = fvad.AudioPipeline()
            ext0 = torch.cuda.ExternalStream(p.cuda_stream(), device=dev)

            def step0():
                p.reset()
                for off in range(0, n0, SR):                       # SimulationInstance.simulateVAD
                    p.push_samples(x0[off:off + SR])
                return len(p.vad_segments)
            step0()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            t0 = time.perf_counter()
            e0.record(ext0)
            for _ in range(reps):
                nseg0 = step0()
            e1.record(ext0)
            torch.cuda.synchronize()
            wall0 = (time.perf_counter() - t0) / reps
            t_dev = e0.elapsed_time(e1) / reps / 1e3
            configs0 = {"workload": "configs[0]: one synthetic 48 kHz mono stream x 60 s, fvad_pipeline_push, 48000-sample pushes from pageable host memory",
                        "value": 60 / t_dev, "unit": UNIT, "ms_per_step": 1e3 * t_dev, "wall_ms_per_step": 1e3 * wall0, "segments": nseg0,
                        "h2d_bytes_per_step": n0 * 4}
            p.close()
            if not args.no_cpu and world == 1:
                c0 = cpu_run(1, 60, faithful=True, threads=1)
                configs0["cpu_single_thread"] = {"value": c0["value"], "unit": UNIT, "wall_s": c0["wall_s"], "kind": c0["kind"],
                                                 "compiler": c0["compiler"], "segments": c0["segments"]}
        except Exception as e:
            errors["configs0"] = repr(e)

    # ---- STFT + band-sum stage alone (configs[2] shape): the bandwidth-bound kernel ------------------
    stft = None
    if not args.no_stft and rank == 0:
        try:
            stft = stft_leg(fvad, torch, dev, peak)
        except Exception as e:
            errors["stft"] = repr(e)

    # ---- many-stream regime (north_star: ">= 1e5 x real-time per B200 on batched 48 kHz streams") --------------------
    # The headline workload has only 20 dependent chains per GPU; this leg shows what the same path does when the batch
    # is wide enough to fill the GPU: every rank, device-resident AND fed from host memory through fvad_batch_push.
    batched = None
    if not args.no_batched:
        try:
            batched = batched_leg(args, fvad, shard, torch, dist, dev, local, rank, world, timed, fill_device, host_rows, host_mem_per_rank)
        except Exception as e:
            errors["batched"] = repr(e)
            barrier()

    # ---- configs[3] as BASELINE.json writes it: 160 streams x 2 h over the N GPUs (strong scaling) -------------------
    strong = None
    if not args.no_strong:
        try:
            strong = strong_leg(args, fvad, shard, torch, dev, local, rank, world, timed, secs)
        except Exception as e:
            errors["configs3_strong"] = repr(e)
            barrier()

    # ---- CPU baseline (rank 0, N = 1 only) --------------------------------------------------------------
    cpu = None
    if not args.no_cpu and rank == 0 and world == 1:
        cs = args.cpu_sample_seconds
        r1 = cpu_run(S, cs, faithful=True)
        r2 = cpu_run(S, max(10, cs // 2), faithful=False)
        cpu = {"value": r1["value"], "unit": UNIT, "cores": min(r1["threads"], r1["host_cpus"] or 1), "threads": r1["threads"],
               "host_cpus": r1["host_cpus"], "kind": r1["kind"], "compiler": r1["compiler"],
               "sample": f"first {cs} s of each of the {S} streams (of {secs} s), one OS thread per stream (src/simulator.zig:217-228), "
                         f"48000-sample pushes, wall {r1['wall_s']:.1f} s",
               "per_stream_wall_s": r1["per_stream_wall_s"], "value_twiddles_hoisted": r2["value"]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "wall_ms_per_step": wall_ms / K, "profiled_step_ms": prof_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic (seeded generator fvad_synth_fill_device, bit-identical on host)",
            "config": {"workload": f"configs[1]: {S} synthetic 48 kHz mono streams x {secs} s per GPU, RNNoise denoiser + VAD, "
                                   f"default plan" + ("" if world == 1 else f"; {world * S} streams sharded by stream (configs[3] at N=8)"),
                       "streams_per_gpu": S, "seconds_per_stream": secs, "fft_size": 2048, "parallelism": f"streams/{world}",
                       "l2": "inputs (%.1f GB per GPU) exceed L2; no flush needed" % (S * n * 4 / 1e9)},
            "clocks": clocks, "e2e": e2e, "e2e_sim": e2e_sim, "gpu_launches": int(launches), "roofline": roofline, "stft": stft,
            "batched": batched, "configs3_strong": strong, "configs0": configs0, "sweep": sweep,
            "cpu_baseline": cpu, "kernel_shares": shares, "segments": n_segments, "segments_sha_first20": sha,
        }
        if errors:
            line["leg_errors"] = errors
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def stft_leg(fvad, torch, dev, peak):
    rows_n, frames_n = 4096, 1406                                # 4096 streams x 60 s per pass (23.6 GB >> L2)
    xs = torch.empty(rows_n, frames_n * 2048, dtype=torch.float32, device=dev)
    fvad.synth_device(xs.data_ptr(), frames_n * 2048, SEED, 1000, rows_n, 0, frames_n * 2048)
    band = torch.empty(rows_n, frames_n, dtype=torch.float32, device=dev)
    f = fvad.FFT(2048, SR)
    cur = torch.cuda.current_stream()

    def run(spectra_ptr, reps):
        for _ in range(3):
            f.stft_batch(xs.data_ptr(), rows_n, frames_n * 2048, frames_n, 2048, 4, 64, band.data_ptr(), spectra_ptr, cur.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            f.stft_batch(xs.data_ptr(), rows_n, frames_n * 2048, frames_n, 2048, 4, 64, band.data_ptr(), spectra_ptr, cur.cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps / 1e3
    t = run(0, 10)
    algo = rows_n * frames_n * 8196
    stft = {"workload": "configs[2] shape: 4096 streams x 60 s per launch, STFT 2048 + speech-band sum (fused, pruned)",
            "kernel": "k_stft2048<false>", "bound": "hbm", "achieved": algo / t / 1e9, "peak": peak, "unit": "GB/s",
            "frac": algo / t / 1e9 / peak, "ms_per_launch": 1e3 * t, "frames_per_s": rows_n * frames_n / t,
            "audio_s_per_s": rows_n * frames_n * 2048 / SR / t, "algorithmic_bytes_per_frame": 8196,
            "algorithmic_bytes_per_launch": algo, "traffic": None}
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        stft["traffic"] = json.load(open(tp)).get("k_stft2048<false>@4096x1406")
    # the spectrum-writing variant (what FFT.fft / Spectrogram.compute return): 12 292 B per frame
    try:
        spec = torch.empty(rows_n * frames_n * 1025, dtype=torch.float32, device=dev)
        t2 = run(spec.data_ptr(), 5)
        algo2 = rows_n * frames_n * 12292
        stft["spectrum_variant"] = {"kernel": "k_stft2048<true>", "bound": "hbm", "achieved": algo2 / t2 / 1e9, "peak": peak, "unit": "GB/s",
                                    "frac": algo2 / t2 / 1e9 / peak, "ms_per_launch": 1e3 * t2, "algorithmic_bytes_per_frame": 12292}
        del spec
    except Exception as e:
        stft["spectrum_variant"] = {"error": repr(e)}
    del xs, band
    f.deinit()
    torch.cuda.empty_cache()
    return stft


def batched_leg(args, fvad, shard, torch, dist, dev, local, rank, world, timed, fill_device, host_rows, host_mem_per_rank):
    sizes = [int(v) for v in str(args.batched_streams).split(",") if v]
    Sb, sb = sizes[0], args.batched_seconds
    nb = sb * SR
    first_id = 5000 + rank * Sb
    xb = torch.empty(Sb, nb, dtype=torch.float32, device=dev)
    fill_device(xb, first_id, Sb, nb)
    torch.cuda.synchronize()
    bb = fvad.PipelineBatch(Sb, device=local)
    extb = torch.cuda.ExternalStream(bb.cuda_stream(), device=dev)

    def result():
        bb.sync()
        per = [[bb.segments(i, 0)] for i in range(Sb)]
        return shard.pack_segments(first_id, per)

    def step_resident():
        bb.reset()
        bb.push_device(xb.data_ptr(), nb, nb, nb)
        return result()
    step_resident()
    reps = 2
    ms, _, rec0 = timed(step_resident, reps, extb)
    tb = ms / reps / 1e3
    out = {"workload": f"{Sb} synthetic 48 kHz mono streams x {sb} s per GPU, full path (RNNoise denoiser + VAD), every rank",
           "streams_per_gpu": Sb, "seconds_per_stream": sb, "n_gpus": world,
           "value": world * Sb * sb / tb, "per_gpu": Sb * sb / tb, "unit": UNIT, "ms_per_step": 1e3 * tb, "segments_rank0": int(len(rec0)),
           "input": "device-resident"}
    # host-fed: the same step through fvad_batch_push, from pinned rows and (N = 1) from pageable rows
    need = Sb * nb * 4
    avail = host_mem_per_rank()
    if avail and need * 1.5 > avail:
        out["e2e"] = {"skipped": f"{need / 1e9:.0f} GB of host rows per rank do not fit beside {avail / 1e9:.0f} GB of free host memory per rank"}
    else:
        hb = torch.empty(Sb, nb, dtype=torch.float32, pin_memory=True)
        hb.copy_(xb)
        torch.cuda.synchronize()

        def step_host(buf):
            def go():
                bb.reset()
                bb.push_rows(host_rows(buf, Sb), nb)
                return result()
            return go
        step_host(hb)()
        ms, wall, rec1 = timed(step_host(hb), reps, extb)
        te = ms / reps / 1e3
        out["e2e"] = {"value": world * Sb * sb / te, "per_gpu": Sb * sb / te, "unit": UNIT, "ms_per_step": 1e3 * te, "wall_ms_per_step": wall / reps,
                      "pinned": True, "h2d_bytes_per_step": need, "h2d_gbs_per_gpu": need / te / 1e9, "d2h_bytes_per_step": int(rec1.nbytes + 8),
                      "vs_resident": tb / te, "segments_equal_device_resident_run": len(rec1) == len(rec0) and bool((rec1 == rec0).all()),
                      "api": "fvad_batch_push, one call per step, rows of one pinned array (2-D copies per tile)"}
        if world == 1 and (not avail or need * 2.5 < avail):
            hp = torch.empty(Sb, nb, dtype=torch.float32)        # pageable: what a Zig allocator hands over
            hp.copy_(hb)
            del hb
            step_host(hp)()
            ms, wall, rec2 = timed(step_host(hp), 1, extb)
            tp = ms / 1e3
            out["e2e_pageable"] = {"value": Sb * sb / tp, "unit": UNIT, "ms_per_step": 1e3 * tp, "wall_ms_per_step": wall, "pinned": False,
                                   "h2d_gbs_per_gpu": need / tp / 1e9, "vs_resident": tb / tp,
                                   "segments_equal_device_resident_run": len(rec2) == len(rec0) and bool((rec2 == rec0).all())}
            del hp
        else:
            del hb
    bb.close()
    del xb
    torch.cuda.empty_cache()
    # stream-count sweep, device-resident, rank 0 at N = 1: the curve between the latency and the throughput regime
    if len(sizes) > 1 and world == 1:
        table = []
        for Sx in sizes[1:]:
            sx = max(60, min(240, (Sb * sb) // Sx // 30 * 30))
            nx = sx * SR
            xt = torch.empty(Sx, nx, dtype=torch.float32, device=dev)
            fill_device(xt, 5000, Sx, nx)
            torch.cuda.synchronize()
            bt = fvad.PipelineBatch(Sx, device=local)
            et = torch.cuda.ExternalStream(bt.cuda_stream(), device=dev)

            def go():
                bt.reset()
                bt.push_device(xt.data_ptr(), nx, nx, nx)
                bt.sync()
            go()
            ms, _, _ = timed(go, 2, et)
            table.append({"streams": Sx, "seconds": sx, "value": Sx * sx / (ms / 2 / 1e3), "ms_per_step": ms / 2})
            bt.close()
            del xt
            torch.cuda.empty_cache()
        out["stream_sweep"] = table
    return out


def strong_leg(args, fvad, shard, torch, dev, local, rank, world, timed, secs):
    """configs[3] as written: 160 streams x 2 h, block-partitioned over the N GPUs.  221 GB of PCM does not fit one GPU, so
    every rank generates its streams on the device in 600-second chunks (two alternating buffers) INSIDE the timed region
    and pushes each chunk with fvad_batch_push_device; at N = 8 this is the weak leg's workload."""
    total = args.strong_streams
    lo, hi = shard.partition(total, world, rank)
    S3 = hi - lo
    chunk = 600 * SR
    n = secs * SR
    bufs = [torch.empty(S3, chunk, dtype=torch.float32, device=dev) for _ in range(2)]
    done = [None, None]
    b3 = fvad.PipelineBatch(S3, device=local)
    ext3 = torch.cuda.ExternalStream(b3.cuda_stream(), device=dev)
    cur = torch.cuda.current_stream()

    def step():
        b3.reset()
        for c, off in enumerate(range(0, n, chunk)):
            m = min(chunk, n - off)
            k = c & 1
            if done[k] is not None:
                done[k].synchronize()                              # the chunk that used this buffer has left the pipeline
            fvad.synth_device(bufs[k].data_ptr(), chunk, SEED, lo, S3, off, m, cur.cuda_stream)
            cur.synchronize()
            b3.push_device(bufs[k].data_ptr(), chunk, chunk, m)
            done[k] = torch.cuda.Event()
            done[k].record(ext3)
        per = [[b3.segments(s, 0)] for s in range(S3)]
        rec = shard.pack_segments(lo, per)
        return shard.gather_segments(rec, device=dev) if world > 1 else rec
    step()
    ms, wall, allrec = timed(step, 1, ext3)
    b3.close()
    del bufs
    torch.cuda.empty_cache()
    t = ms / 1e3
    return {"workload": f"configs[3]: {total} synthetic 48 kHz mono streams x {secs} s sharded by stream over {world} GPU(s) "
                        f"({S3} per GPU), generated on the device in 600 s chunks inside the timed region",
            "scaling": "strong", "n_gpus": world, "streams_per_gpu": S3, "value": total * secs / t, "per_gpu": total * secs / t / world,
            "unit": UNIT, "ms_per_step": ms, "wall_ms_per_step": wall,
            "segments": None if allrec is None else int(len(allrec)), "segments_sha_first20": sha_first(allrec),
            "limiter": "streams per GPU: below ~110 chains per GPU the per-stream sequential chains (biquad first) set the pace, "
                       "not SM throughput and not NCCL (DESIGN.md section 5)"}


SWEEP_KEYS = ("speech_threshold_factor", "short_term_speech_avg_sec", "min_consecutive_sec_to_open", "max_speech_gap_sec", "min_vad_duration_sec")


def sweep_grid():
    """256 VADMachine parameter sets: 8 threshold factors x 4 short-term windows x 2 x 2 x 2."""
    import itertools
    f = [6.0, 9.0, 12.0, 15.0, 18.0, 24.0, 30.0, 40.0]
    return [dict(zip(SWEEP_KEYS, g)) for g in itertools.product(f, (0.1, 0.2, 0.3, 0.5), (0.1, 0.2), (0.5, 2.0), (0.3, 0.7))]


def sweep_leg(args, fvad, b, S, secs, first_stream, torch, dev, do_cpu):
    """BASELINE.json configs[4]: 256 parameter sets x the batch's streams, scanned on the device over the features the run
    left in HBM (fvad_batch_sweep), lists kept on the device and scored there (fvad_eval_stats_batch_device)."""
    import numpy as np
    from fvad import evaluator as ev
    grid = sweep_grid()
    cfgs = [fvad.default_machine_config(**g) for g in grid]
    K, max_segs = len(cfgs), 1024
    n_rec = b.feature_count()
    d_segs = torch.empty(S * K * max_segs * fvad.SEG_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    d_counts = torch.zeros(S * K, dtype=torch.int32, device=dev)
    ext = torch.cuda.ExternalStream(b.cuda_stream(), device=dev)
    torch.cuda.synchronize()

    def go():
        b.sweep(cfgs, max_segs=max_segs, d_segs=d_segs.data_ptr(), d_counts=d_counts.data_ptr())
    go()
    reps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(ext)
    for _ in range(reps):
        go()
    e1.record(ext)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / reps
    t = e0.elapsed_time(e1) / reps / 1e3
    counts = d_counts.cpu().numpy().view(np.uint32)
    # the consumer: Evaluator statistics of every (stream, config) list against per-stream labels (the default plan's own
    # segments serve as labels), the lists never leaving the device
    refs = []
    for s_ in range(S):
        seg = b.segments(s_, 0)
        r = np.zeros(len(seg), ev.REF_DTYPE)
        r["from_sec"] = seg["sample_from"].astype(np.float32) / np.float32(SR)
        r["to_sec"] = seg["sample_to"].astype(np.float32) / np.float32(SR)
        refs.append(r)
    set_ref = np.repeat(np.arange(S, dtype=np.uint32), K)
    t1 = time.perf_counter()
    st = ev.stats_batch_device(d_segs.data_ptr(), S * K, max_segs, d_counts.data_ptr(), set_ref, refs, SR, ev.StatConfig(0.7, 5, 10, 5))
    t_stats = time.perf_counter() - t1
    k_def = next(i for i, g in enumerate(grid) if g == dict(zip(SWEEP_KEYS, (18.0, 0.2, 0.2, 2.0, 0.7))))
    out = {"workload": f"configs[4]: {K} VADMachine parameter sets x {S} streams x {n_rec} STFT records ({secs} s), shared features kept in HBM",
           "api": "fvad_batch_sweep (device lists) + fvad_eval_stats_batch_device", "scans": S * K, "records_per_scan": n_rec,
           "ms_per_sweep": 1e3 * t, "wall_ms_per_sweep": 1e3 * wall, "scans_per_s": S * K / t, "record_steps_per_s": S * K * n_rec / t,
           "segments_total": int(np.minimum(counts, max_segs).sum()), "lists_truncated": int((counts > max_segs).sum()),
           "stats_ms": 1e3 * t_stats, "default_config_f_score_vs_own_labels": float(st[k_def]["f_score"]),
           "hbm_bytes_read": S * K * n_rec * 12, "note": "one thread per scan, state in registers; the binding limit is the per-record "
                                                          "dependent f64 ring sums (latency), not bandwidth"}
    if do_cpu:
        # cpu_baseline of THIS workload: the oracle's VADMachine (restated VADMachine.zig / RollingAverage.zig: the whole long ring is
        # re-summed per record) on a bounded sample: one scan per host thread, first `n_cpu` records; lists compared with the device's
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        from concurrent.futures import ThreadPoolExecutor
        import oracle
        load, kind, compiler = cpu_kind()
        orc = oracle.Oracle(load if load != "reference_fast" else "reference")     # lists are COMPARED: the parity build
        threads = os.cpu_count() or 1
        n_cpu = min(n_rec, 12000)
        recs = [b.features(s_)[:n_cpu] for s_ in range(min(S, 4))]
        picks = [(i % len(recs), (37 * i) % K) for i in range(threads)]

        def one(sk):
            s_, k = sk
            return orc.vadm_scan(orc.default_config(**grid[k]), recs[s_])
        with ThreadPoolExecutor(threads) as ex:
            t2 = time.perf_counter()
            lists = list(ex.map(one, picks))
            t_cpu = time.perf_counter() - t2
        # the device's lists for the same scans, restricted to what the first n_cpu records can have decided
        segs_h = d_segs.cpu().numpy().view(fvad.SEG_DTYPE).reshape(S * K, max_segs)
        equal = True
        for (s_, k), want in zip(picks, lists):
            got = segs_h[s_ * K + k, :min(counts[s_ * K + k], max_segs)]
            equal &= got[:len(want)].tobytes() == want.tobytes()
        out["cpu_baseline"] = {"record_steps_per_s": len(picks) * n_cpu / t_cpu, "unit": "record steps/s", "cores": threads, "kind": kind,
                               "sample": f"{len(picks)} scans x first {n_cpu} records, one scan per thread, wall {t_cpu:.1f} s",
                               "full_sweep_equivalent_s": S * K * n_rec / (len(picks) * n_cpu / t_cpu),
                               "lists_equal_device_prefix": bool(equal)}
    return out


def main():
    # Libraries (NCCL's version banner, torchrun notices) write to stdout; the contract is ONE JSON line there.
    # Point fd 1 at stderr for the whole run and print the JSON line through a saved copy of the real stdout.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
