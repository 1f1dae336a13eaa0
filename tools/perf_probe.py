"""Quick timing probe (not the bench): S streams x secs through the full pipeline, device-resident input."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
import torch, fvad
S = int(sys.argv[1]) if len(sys.argv) > 1 else 20
secs = int(sys.argv[2]) if len(sys.argv) > 2 else 120
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
use_dn = int(sys.argv[4]) if len(sys.argv) > 4 else 1
n = secs * 48000
x = torch.empty(S, n, dtype=torch.float32, device="cuda")
fvad.synth_device(x.data_ptr(), n, 0xF1AD5EED00000000, 0, S, 0, n)
torch.cuda.synchronize()
b = fvad.PipelineBatch(S, use_denoiser=bool(use_dn))
for r in range(reps):
    b.reset()
    t0 = time.time()
    b.push_device(x.data_ptr(), n, n, n)
    b.sync()
    dt = time.time() - t0
    nseg = sum(len(b.segments(s)) for s in range(S))
    print(f"rep {r}: {dt*1e3:.1f} ms  -> {S*secs/dt:.0f} x real-time, segments {nseg}", flush=True)
