"""Turn the files tools/gpu_profile.sh leaves in gpurun_out/ into the summaries committed under profiles/:
  <tag>_ncu_launch_summary.csv   per-kernel launches / total ms / share from the ncu launch list
  <tag>_ncu_full_metrics.txt     the metrics that matter from each `ncu --set full` capture
  ncu_traffic.json               dram bytes (read + write) per launch, read by bench.py for roofline.traffic
usage: python tools/ncu_summarise.py r01e [gpurun_out]"""
import csv, io, json, os, re, subprocess, sys

tag = sys.argv[1]
src = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(ROOT, "profiles")

KEEP = re.compile(r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|dram__throughput\.avg\.pct|launch__(grid_size|block_size|registers_per_thread$|"
                  r"occupancy_limit|shared_mem_per_block)|sm__warps_active\.avg\.pct|smsp__issue_active\.avg\.pct|sm__throughput\.avg\.pct|"
                  r"l1tex__throughput\.avg\.pct|l1tex__data_pipe_lsu_wavefronts\.avg\.pct|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|"
                  r"lts__throughput\.avg\.pct|sm__inst_executed_pipe_(fma|alu|lsu|fp64|xu)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__pipe_(fma|fmaheavy|fp64|tensor)_cycles_active\.avg\.pct_of_peak_sustained_elapsed|smsp__inst_executed\.sum$|"
                  r"smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio|sm__sass_thread_inst_executed_op_f(add|mul|fma)_pred_on\.sum$)")


def short(name):
    m = re.match(r"(?:void )?(?:fvad::)?(?:rn::)?(\w+(?:<[^>]*>)?)", name.strip())
    return m.group(1) if m else name


# ---- launch list --------------------------------------------------------------------------------
p = os.path.join(src, "launches.csv")
if os.path.exists(p):
    lines = [ln for ln in open(p) if ln.startswith('"')]
    rows = list(csv.reader(io.StringIO("".join(lines))))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg, total = {}, 0.0
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui] in ("us", "usecond") else v
        k = short(r[ki])
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v; total += v
    with open(os.path.join(out, f"{tag}_ncu_launch_summary.csv"), "w") as f:
        f.write("# ncu launch list summary (gpu__time_duration.sum, --clock-control none); command: see tools/gpu_profile.sh\n")
        f.write(f"# {len(rows) - 1} launches captured, total {total:.1f} ms; per-launch times are cold-cache and serialised: compare SHARES\n")
        f.write("kernel,launches,total_ms,share\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{n},{t:.3f},{t / total:.4f}\n")
    print("wrote launch summary:", len(rows) - 1, "launches")

# ---- full captures -----------------------------------------------------------------------------
traffic_path = os.path.join(out, "ncu_traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
with open(os.path.join(out, f"{tag}_ncu_full_metrics.txt"), "w") as f:
    for rep in sorted(x for x in os.listdir(src) if x.endswith(".ncu-rep")):
        r = subprocess.run(["ncu", "-i", os.path.join(src, rep), "--page", "raw", "--csv"], capture_output=True, text=True)
        rows = list(csv.reader(io.StringIO(r.stdout)))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        for row in rows[2:]:
            name = row[hdr.index("Kernel Name")]
            f.write(f"== {rep}: {name.strip()}\n")
            vals = {}
            for h, u, v in zip(hdr, units, row):
                vals[h] = (v, u)
                if KEEP.match(h):
                    f.write(f"{h} = {v} {u}\n")
            try:
                def nbytes(key):
                    v, u = vals[key]
                    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
                k = short(name)
                if k.startswith("k_stft2048"):
                    k = "k_stft2048<true>" if re.match(r"k_stft2048<(?:\(bool\))?(1|true)", k) else "k_stft2048<false>"
                else:
                    k = re.sub(r"<.*", "", k)              # the library's profile names carry no template arguments
                    if k in ("k_rn_biquad_chain", "k_rn_biquad_chain_s"):
                        k = "k_rn_biquad"                  # the name the library's profile (and bench.py) uses for the chain kernel
                tb = nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")
                if k.startswith("k_stft2048") and tb > 1e10:
                    k += "@4096x1406"                      # the stft leg's launch (configs[2] chunk), not a pipeline tile
                if rep.startswith("prof_"):                # t_*: throughput-regime captures (592 streams), not the bench's launch shapes
                    traffic[k] = tb
            except (KeyError, ValueError):
                pass
            f.write("\n")
json.dump(traffic, open(traffic_path, "w"), indent=1, sort_keys=True)
print("traffic:", traffic)
