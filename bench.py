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

`value`  : inputs resident in HBM before the timed region (27.6 GB per GPU >> L2, so no L2 flush is needed).
`e2e`    : the same step through the reference-facing call with HOST buffers (fvad_batch_push = AudioPipeline.
           pushSamples): H2D copies from pinned memory and the D2H segment read are inside the timed region.
`roofline`: the kernel with the largest share of the step, timed live with CUDA events on its own stream
           (the library's per-launch event profiler) against the measured HBM peak of MEASURED_PEAKS.json.
`stft`   : BASELINE.json configs[2] shape (4096 streams, STFT + band sum only), the bandwidth-bound stage.
`cpu_baseline` / `--impl reference`: the reference's own C sources (oracle/_ref) under the restated Zig glue
           on the box's host cores, one thread per stream as src/simulator.zig:217-228 does.
"""
import argparse
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
    ap.add_argument("--no-stft", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the many-stream leg")
    ap.add_argument("--batched-streams", type=int, default=1184)
    ap.add_argument("--batched-seconds", type=int, default=150)
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's C libraries + restated Zig glue, one thread per stream
# ------------------------------------------------------------------------------------------------
def cpu_run(n_streams, seconds, faithful=True, first_stream=0):
    """Returns (audio_s_per_s, wall_s, threads, kind, total_segments)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from concurrent.futures import ThreadPoolExecutor
    import fvad
    from oracle import Oracle, have_reference
    kind = "reference" if have_reference() else "port"
    orc = Oracle(kind)
    threads = max(1, min(n_streams, os.cpu_count() or 1))
    n = seconds * SR
    with ThreadPoolExecutor(threads) as ex:           # input generation is outside the timed region
        pcm = list(ex.map(lambda s: fvad.synth_host(SEED, first_stream + s, 0, n), range(n_streams)))

        def one(s):                                    # SimulationInstance.simulateVAD: 48000-sample pushes
            r = orc.run_pipeline([pcm[s]], faithful=faithful, push=48000)
            return len(r["segments"][0])
        t0 = time.perf_counter()
        segs = list(ex.map(one, range(n_streams)))
        wall = time.perf_counter() - t0
    return n_streams * seconds / wall, wall, threads, kind, int(sum(segs))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # the WHOLE job's streams (weak scaling: streams-per-GPU x N), on the host cores of the box, by rank 0 alone
    S, secs = args.streams * max(1, args.gpus), args.ref_sample_seconds
    for _ in range(args.warmup):
        cpu_run(S, max(10, secs // 8))
    walls, segs = [], 0
    for _ in range(args.steps):
        v, wall, threads, kind, segs = cpu_run(S, secs)
        walls.append(wall)
    wall = sum(walls) / len(walls)
    value = S * secs / wall
    sample = (f"first {secs} s of each of the {S} streams per step (of 7200 s), one thread per stream as "
              f"src/simulator.zig:217-228, {threads} threads, KissFFT state regenerated per FFT as src/FFT.zig:84")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (seeded generator, same streams as the b200 arm)",
        "config": {"workload": f"configs[1]: {args.streams} synthetic 48 kHz mono streams x 2 h per GPU, RNNoise denoiser + VAD, "
                               f"default plan" + ("" if args.gpus <= 1 else f"; {S} streams, the {args.gpus}-GPU job's whole workload (configs[3] at N=8)"),
                   "streams": S, "streams_per_gpu": args.streams, "sample_seconds_per_stream": secs, "push": 48000},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
                         "compiler": "gcc -O2 -ffp-contract=off (zig ReleaseFast unavailable on this box)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "segments": segs,
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
        dist.init_process_group("nccl", device_id=dev)
    S, secs = args.streams, args.seconds
    n = secs * SR
    first_stream = rank * S                      # weak scaling: every rank owns its own S streams
    K, W = args.steps, args.warmup

    # ---- inputs resident in HBM ---------------------------------------------------------------
    x = torch.empty(S, n, dtype=torch.float32, device=dev)
    CH = 600 * SR
    for off in range(0, n, CH):
        m = min(CH, n - off)
        fvad.synth_device(x.data_ptr() + 4 * off, n, SEED, first_stream, S, off, m)
    torch.cuda.synchronize()
    b = fvad.PipelineBatch(S, device=local)
    ext = torch.cuda.ExternalStream(b.cuda_stream(), device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def collect():
        per = [[b.segments(s, 0)] for s in range(S)]
        rec = shard.pack_segments(first_stream, per)
        allrec = shard.gather_segments(rec, device=dev) if world > 1 else rec
        return rec, allrec

    def step_device():
        b.reset()
        b.push_device(x.data_ptr(), n, n, n)
        return collect()

    def timed(fn, k):
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

    for _ in range(W):
        step_device()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = fvad.launch_count()
    dev_ms, wall_ms, (rec, allrec) = timed(step_device, K)
    launches = fvad.launch_count() - launches0
    # one more step with the library's per-launch CUDA events on (two events around every launch, on the launch's
    # own stream): the per-kernel durations the roofline uses.  Kept out of `value` because the extra event
    # records sit between the kernels of the four stage streams.
    fvad.profile_enable(True)
    prof_ms, _, _ = timed(step_device, 1)
    fvad.profile_enable(False)
    prof = fvad.profile_report()
    clocks = sampler.stop()
    ms_per_step = dev_ms / K
    value = world * S * secs / (ms_per_step / 1e3)
    n_segments = int(len(allrec)) if allrec is not None else None

    # ---- roofline of the dominant kernel ---------------------------------------------------------
    peak, peak_src = peaks()
    roofline = None
    shares = {}
    if prof:
        tot = sum(v["total_ms"] for v in prof.values())
        shares = {k: round(v["total_ms"] / tot, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"])}
        top = max(prof.items(), key=lambda kv: kv[1]["total_ms"])
        name, info = top
        frames = S * (n // 480)                      # the profiled step
        stft_frames = S * (n // 2048)
        per, unit = ALGO_BYTES.get(name, (None, None))
        traffic = None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(name)
        if per is not None:
            units = frames if unit == "frame" else stft_frames
            algo = per * units / info["launches"]                    # bytes per launch
            dur_s = info["total_ms"] / 1e3 / info["launches"]        # average launch duration
            ach = algo / dur_s / 1e9
            roofline = {"kernel": name, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                        "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": algo,
                        "avg_launch_ms": 1e3 * dur_s, "launches": info["launches"], "share_of_step": shares[name],
                        "note": "latency-bound sequential scan: the binding limit is the dependent chain per stream, not HBM "
                                "(DESIGN.md); the bandwidth-bound stage is reported under 'stft'"}
        else:
            roofline = {"kernel": name, "bound": "hbm", "achieved": None, "peak": peak, "unit": "GB/s", "frac": None,
                        "traffic": traffic, "share_of_step": shares[name]}

    # ---- e2e: host buffers through fvad_batch_push --------------------------------------------------
    e2e = None
    if not args.no_e2e:
        need = S * n * 4
        avail = 0
        try:
            for ln in open("/proc/meminfo"):
                if ln.startswith("MemAvailable"):
                    avail = int(ln.split()[1]) * 1024
        except OSError:
            pass
        win = n
        if avail and need * world * 2 > avail:                    # keep well inside host RAM: bounded window, pushed repeatedly
            win = max(600 * SR, int(avail / (world * 8) / (S * 4)) // (600 * SR) * (600 * SR))
            win = min(win, n)
        h = torch.empty(S, win, dtype=torch.float32, pin_memory=True)
        for off in range(0, win, CH):
            m = min(CH, win - off)
            h[:, off:off + m].copy_(x[:, off:off + m])
        torch.cuda.synchronize()
        base = h.data_ptr()
        rows = [base + 4 * win * r for r in range(S)]

        def step_host():
            b.reset()
            done = 0
            while done < n:
                m = min(win, n - done)
                b.push_rows(rows, m)
                done += m
            return collect()
        step_host()
        e_ms, e_wall, (erec, _) = timed(step_host, K)
        e2e = {"value": world * S * secs / (e_ms / K / 1e3), "unit": UNIT, "h2d_bytes_per_step": S * n * 4,
               "d2h_bytes_per_step": int(erec.nbytes + 8), "ms_per_step": e_ms / K, "wall_ms_per_step": e_wall / K,
               "host_window_s": win // SR, "pinned": True,
               "api": "fvad_batch_push (AudioPipeline.pushSamples drop-in), host PCM in, segment list out"}
        if win == n:
            same = len(erec) == len(rec) and bool((erec == rec).all())
            e2e["segments_equal_device_resident_run"] = same
        del h
    del x
    torch.cuda.empty_cache()

    # ---- STFT + band-sum stage alone (configs[2] shape): the bandwidth-bound kernel ------------------
    stft = None
    if not args.no_stft and rank == 0:
        rows_n, frames_n = 4096, 1406                                # 4096 streams x 60 s per pass (23.6 GB >> L2)
        xs = torch.empty(rows_n, frames_n * 2048, dtype=torch.float32, device=dev)
        fvad.synth_device(xs.data_ptr(), frames_n * 2048, SEED, 1000, rows_n, 0, frames_n * 2048)
        band = torch.empty(rows_n, frames_n, dtype=torch.float32, device=dev)
        f = fvad.FFT(2048, SR)
        cur = torch.cuda.current_stream()
        for _ in range(3):
            f.stft_batch(xs.data_ptr(), rows_n, frames_n * 2048, frames_n, 2048, 4, 64, band.data_ptr(), 0, cur.cuda_stream)
        torch.cuda.synchronize()
        reps = 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            f.stft_batch(xs.data_ptr(), rows_n, frames_n * 2048, frames_n, 2048, 4, 64, band.data_ptr(), 0, cur.cuda_stream)
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / reps / 1e3
        algo = rows_n * frames_n * 8196
        stft = {"workload": "configs[2] shape: 4096 streams x 60 s per launch, STFT 2048 + speech-band sum (fused, pruned)",
                "kernel": "k_stft2048<false>", "bound": "hbm", "achieved": algo / t / 1e9, "peak": peak, "unit": "GB/s",
                "frac": algo / t / 1e9 / peak, "ms_per_launch": 1e3 * t, "frames_per_s": rows_n * frames_n / t,
                "audio_s_per_s": rows_n * frames_n * 2048 / SR / t, "algorithmic_bytes_per_frame": 8196,
                "algorithmic_bytes_per_launch": algo, "traffic": None}
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):
            stft["traffic"] = json.load(open(tp)).get("k_stft2048<false>@4096x1406")
        launches += 0
        del xs, band
        f.deinit()

    # ---- many-stream regime (north_star: ">= 1e5 x real-time per B200 on batched 48 kHz streams") --------------------
    # The headline workload has only 20 dependent chains per GPU; this leg shows what the same path does when the batch
    # is wide enough to fill the GPU.  Device-resident input, same API, rank 0 at N = 1 only.
    batched = None
    if not args.no_batched and rank == 0 and world == 1:
        Sb, sb = args.batched_streams, args.batched_seconds
        nb = sb * SR
        xb = torch.empty(Sb, nb, dtype=torch.float32, device=dev)
        for off in range(0, nb, CH):
            m = min(CH, nb - off)
            fvad.synth_device(xb.data_ptr() + 4 * off, nb, SEED, 5000, Sb, off, m)
        torch.cuda.synchronize()
        bb = fvad.PipelineBatch(Sb, device=local)
        extb = torch.cuda.ExternalStream(bb.cuda_stream(), device=dev)

        def step_batched():
            bb.reset()
            bb.push_device(xb.data_ptr(), nb, nb, nb)
            bb.sync()
            return sum(len(bb.segments(i)) for i in range(Sb))
        step_batched()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 2
        e0.record(extb)
        for _ in range(reps):
            nseg_b = step_batched()
        e1.record(extb)
        torch.cuda.synchronize()
        tb = e0.elapsed_time(e1) / reps / 1e3
        batched = {"workload": f"{Sb} synthetic 48 kHz mono streams x {sb} s, full path (RNNoise denoiser + VAD), device-resident",
                   "value": Sb * sb / tb, "unit": UNIT, "ms_per_step": 1e3 * tb, "segments": nseg_b}
        bb.close()
        del xb
        torch.cuda.empty_cache()

    # ---- CPU baseline (rank 0, N = 1 only) --------------------------------------------------------------
    cpu = None
    if not args.no_cpu and rank == 0 and world == 1:
        cs = args.cpu_sample_seconds
        v, wall, threads, kind, _ = cpu_run(S, cs, faithful=True)
        v2, wall2, _, _, _ = cpu_run(S, cs, faithful=False)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
               "sample": f"first {cs} s of each of the {S} streams, one thread per stream (src/simulator.zig:217-228), "
                         f"wall {wall:.1f} s; gcc -O2 -ffp-contract=off stand-in for zig ReleaseFast",
               "value_twiddles_hoisted": v2, "host_cpus": os.cpu_count()}

    b.close()
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "wall_ms_per_step": wall_ms / K, "profiled_step_ms": prof_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic (seeded generator fvad_synth_fill_device, bit-identical on host)",
            "config": {"workload": f"configs[1]: {S} synthetic 48 kHz mono streams x {secs} s per GPU, RNNoise denoiser + VAD, "
                                   f"default plan" + ("" if world == 1 else f"; {world * S} streams sharded by stream (configs[3] at N=8)"),
                       "streams_per_gpu": S, "seconds_per_stream": secs, "fft_size": 2048, "parallelism": f"streams/{world}",
                       "l2": "inputs (%.1f GB per GPU) exceed L2; no flush needed" % (S * n * 4 / 1e9)},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "stft": stft,
            "batched": batched,
            "cpu_baseline": cpu, "kernel_shares": shares, "segments": n_segments,
        }
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
