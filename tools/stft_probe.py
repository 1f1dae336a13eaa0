"""STFT + band-sum stage alone (configs[2] shape): GB/s of algorithmic bytes. usage: stft_probe.py [rows=4096] [frames=1406]"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
import torch, fvad
for k, v in os.environ.items():
    if k.startswith("FVADK_"):
        fvad.debug_set(k[6:].lower(), v)
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 1406
xs = torch.empty(rows, frames * 2048, dtype=torch.float32, device="cuda")
fvad.synth_device(xs.data_ptr(), frames * 2048, 0xF1AD5EED00000000, 1000, rows, 0, frames * 2048)
band = torch.empty(rows, frames, dtype=torch.float32, device="cuda")
spec = None
f = fvad.FFT(2048, 48000)
cur = torch.cuda.current_stream()
for full in (0, 1):
    if full:
        frames2 = frames // 4
        spec = torch.empty(rows, frames2, 1025, dtype=torch.float32, device="cuda")
    fr = frames2 if full else frames
    call = lambda: f.stft_batch(xs.data_ptr(), rows, frames * 2048, fr, 2048, 4, 64, band.data_ptr(), spec.data_ptr() if full else 0, cur.cuda_stream)
    for _ in range(3): call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): call()
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 10 / 1e3
    per = 12292 if full else 8196
    print(f"{'spectrum' if full else 'band-sum'} variant: {rows}x{fr} frames {t*1e3:.2f} ms -> {rows*fr*per/t/1e9:.0f} GB/s algorithmic ({rows*fr*per/t/1e9/6521.4*100:.1f}% of 6521 GB/s), {rows*fr/t/1e6:.0f} Mframes/s")
