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
exit $rc
