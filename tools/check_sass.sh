#!/usr/bin/env bash
# Exactness guard on the built library's SASS (run after `make -C formula-vad_b200/csrc`):
#  * every FFMA2 of the STFT kernels must be a product-forming fma(a, b, -0) (stft.cu: prod2): its addend is the
#    one register (or uniform register) that holds StftTwA::neg_zero.  A multiply that ptxas contracted into a
#    following add would show up as an FFMA2 with some other addend and would drop a rounding KissFFT performs.
#  * the STFT kernels must use the packed f32x2 pipe (FADD2 present) and the staged variant the TMA engine (UBLKCP).
set -euo pipefail
LIB="${1:-$(dirname "$0")/../formula-vad_b200/libfvad_cuda.so}"
SASS="$(mktemp)"; trap 'rm -f "$SASS"' EXIT
cuobjdump -sass "$LIB" > "$SASS"
rc=0
for v in ILb0ELb0ELb0 ILb0ELb1ELb0 ILb1ELb0ELb0 ILb1ELb0ELb1 ILb1ELb1ELb1; do
    body="$(awk -v v="$v" '$0 ~ "Function : .*k_stft2048" v {p=1; next} p && /Function : / {p=0} p' "$SASS")"
    [ -n "$body" ] || { echo "check_sass: k_stft2048<$v> not found"; rc=1; continue; }
    addends="$(grep FFMA2 <<<"$body" | sed 's/.*FFMA2 [^,]*,[^,]*,[^,]*, *\([A-Z0-9]*\).*/\1/' | sort -u)"
    n_vec="$(grep -c '^R' <<<"$addends" || true)"; n_uni="$(grep -c '^UR' <<<"$addends" || true)"
    n_fadd2="$(grep -c FADD2 <<<"$body" || true)"
    echo "k_stft2048<$v>: FFMA2 addends: $(tr '\n' ' ' <<<"$addends") FADD2=$n_fadd2"
    if [ "$n_vec" -gt 1 ] || [ "$n_uni" -gt 1 ]; then echo "  FAIL: contracted FFMA2 suspected"; rc=1; fi
    if [ "$n_fadd2" -lt 100 ]; then echo "  FAIL: packed f32x2 arithmetic missing"; rc=1; fi
done
for v in ILb0ELb1ELb0 ILb1ELb1ELb1; do
    n_tma="$(awk -v v="$v" '$0 ~ "Function : .*k_stft2048" v {p=1; next} p && /Function : / {p=0} p' "$SASS" | grep -c UBLKCP || true)"
    echo "k_stft2048<$v>: UBLKCP=$n_tma"
    [ "$n_tma" -ge 2 ] || { echo "  FAIL: bulk copies missing"; rc=1; }
done
# The biquad chain lane of the headline shape (k_rn_biquad_chain_s1<48>): its fast path must close the loop itself -- a
# uniform backward branch straight from the end of the 48-sample body, no double-precision instruction inside -- and the
# static schedule of that body (sum of the stall counts in the control words) must stay at the measured 31.7 cycles per
# sample.  ptxas laid the same source out differently inside another kernel (two taken branches per iteration: 40.3 instead
# of 36.8 cycles per sample on the GPU); this catches such a change without one.
python3 - "$SASS" <<'PY' || rc=1
import re, sys
lines = open(sys.argv[1]).read().split("\n")
bad = 0
for key, label in (("k_rn_biquad_chain_s1ILi48E", "k_rn_biquad_chain_s1<48>"), ("k_rn_biquad_chain_sILi48ELi2E", "k_rn_biquad_chain_s<48, 2>"),
                   ("k_rn_biquad_chain_sILi48ELi4E", "k_rn_biquad_chain_s<48, 4>")):
    start = next((i for i, l in enumerate(lines) if "Function : " in l and key in l), None)
    if start is None:
        print("check_sass: %s not found" % label); bad = 1; continue
    ins, i = [], start + 1
    while i < len(lines) and "Function : " not in lines[i]:
        m = re.match(r"\s*/\*([0-9a-f]{4})\*/\s+(.*?)\s*/\* (0x[0-9a-f]{16}) \*/", lines[i])
        if m and i + 1 < len(lines):
            m2 = re.match(r"\s*/\* (0x[0-9a-f]{16}) \*/", lines[i + 1])
            ins.append((int(m.group(1), 16), m.group(2), (int(m2.group(1), 16) >> 41) & 0xf)); i += 2
        else:
            i += 1
    best = None
    for a, tx, st in ins:
        m = re.search(r"BRA\.U\s+!?UP\d,\s*0x([0-9a-f]+)", tx)
        if m and int(m.group(1), 16) < a:
            body = [x for x in ins if int(m.group(1), 16) <= x[0] <= a]
            if not any("DFMA" in x[1] or "DADD" in x[1] for x in body) and (best is None or len(body) > len(best)):
                best = body
    if best is None:
        print("%s: FAIL: the fast path has no back-edge of its own" % label); bad = 1; continue
    per, cyc = len(best) / 48.0, sum(x[2] for x in best) / 48.0
    print("%s: fast loop %d instructions (%.2f per sample), static schedule %.2f cycles per sample" % (label, len(best), per, cyc))
    if per > 22.5 or cyc > 33.0:
        print("  FAIL: the lane's loop changed (measured code: 21.6-21.8 instructions, 31.7-32.2 cycles per sample)"); bad = 1
sys.exit(bad)
PY
exit $rc
