"""Quick timing probe (not the bench): S streams x secs through the full pipeline, device-resident input.
usage: perf_probe.py [S=20] [secs=120] [reps=2] [use_denoiser=1] [profile=0]"""
import sys, os, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
import torch, fvad
S = int(sys.argv[1]) if len(sys.argv) > 1 else 20
secs = int(sys.argv[2]) if len(sys.argv) > 2 else 120
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
use_dn = int(sys.argv[4]) if len(sys.argv) > 4 else 1
prof = int(sys.argv[5]) if len(sys.argv) > 5 else 0
host = int(sys.argv[6]) if len(sys.argv) > 6 else 0
for k, v in os.environ.items():          # FVADK_<knob>=<value> -> fvad_debug_set (the library itself reads no environment)
    if k.startswith("FVADK_"):
        fvad.debug_set(k[6:].lower(), v)
n = secs * 48000
x = torch.empty(S, n, dtype=torch.float32, device="cuda")
fvad.synth_device(x.data_ptr(), n, 0xF1AD5EED00000000, 0, S, 0, n)
if os.environ.get("PROBE_INPUT") == "noise":      # white noise of constant level instead of the synthetic streams
    x.uniform_(-0.06, 0.06)
elif os.environ.get("PROBE_INPUT") == "tone":
    t = torch.arange(n, device="cuda", dtype=torch.float32)
    x.copy_((0.1 * torch.sin(t * (2 * 3.141592653589793 * 220.0 / 48000.0))).expand(S, n))
torch.cuda.synchronize()
b = fvad.PipelineBatch(S, use_denoiser=bool(use_dn))
if host:
    h = torch.empty(S, n, dtype=torch.float32, pin_memory=True); h.copy_(x); torch.cuda.synchronize()
    rows = [h.data_ptr() + 4 * n * r for r in range(S)]
for r in range(reps):
    b.reset()
    if prof and r == reps - 1:
        fvad.profile_enable(True)
    t0 = time.time()
    if host:
        b.push_rows(rows, n)
    else:
        b.push_device(x.data_ptr(), n, n, n)
    b.sync()
    dt = time.time() - t0
    nseg = sum(len(b.segments(s)) for s in range(S))
    print(f"S={S} secs={secs} dn={use_dn} host={host} rep {r}: {dt*1e3:.1f} ms  -> {S*secs/dt:.0f} x real-time, segments {nseg}", flush=True)
if prof:
    rep = fvad.profile_report()
    tot = sum(v["total_ms"] for v in rep.values())
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1]["total_ms"]):
        print(f"  {k:24s} {v['launches']:6d} launches {v['total_ms']:10.2f} ms {100*v['total_ms']/tot:5.1f}%")
