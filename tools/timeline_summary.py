"""Summarise a FVAD_TIMELINE csv (kernel,start_ms,end_ms): per kernel busy time, and how the wall time splits into
slices by WHICH classes of kernels were in flight (frame-parallel / RNN / latency-bound scans).
usage: python tools/timeline_summary.py gpurun_out/timeline_592.csv"""
import csv, sys
rows = [(r["kernel"], float(r["start_ms"]), float(r["end_ms"])) for r in csv.DictReader(open(sys.argv[1]))]
LAT = {"k_rn_biquad", "k_vadm", "k_rn_pitch_seq", "k_rn_gain_scan", "k_rn_compact"}
def cls(k):
    return "L" if k in LAT else "R" if k == "k_rn_rnn" else "F"
t0 = min(r[1] for r in rows); t1 = max(r[2] for r in rows)
ev = []
for k, a, b in rows:
    ev.append((a, 1, cls(k))); ev.append((b, -1, cls(k)))
ev.sort()
cnt = {"L": 0, "R": 0, "F": 0}; last = t0; acc = {}
for t, d, c in ev:
    key = "".join(x for x in "FRL" if cnt[x] > 0) or "idle"
    acc[key] = acc.get(key, 0.0) + (t - last); last = t
    cnt[c] += d
print(f"wall {t1 - t0:.1f} ms, {len(rows)} launches")
for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
    print(f"  in flight {k:5s} {v:9.1f} ms {100 * v / (t1 - t0):5.1f}%")
busy = {}
for k, a, b in rows:
    busy[k] = busy.get(k, 0.0) + (b - a)
for k, v in sorted(busy.items(), key=lambda kv: -kv[1])[:12]:
    print(f"  {k:22s} sum of launch spans {v:9.1f} ms")
