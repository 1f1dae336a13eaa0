#!/bin/bash
mkdir -p gpurun_out
{
for P in "9,9,9,9,9,9" "0,9,9,9,9,0" "0,5,4,3,2,1" "0,3,2,1,1,0" "0,2,2,1,9,2" "0,5,5,5,5,1" "0,4,3,2,1,0"; do
  for SL in 3 5; do
    echo "== FVAD_PRIO=$P FVAD_SLOTS=$SL"; FVAD_PRIO=$P FVAD_SLOTS=$SL python tools/perf_probe.py 592 120 2 1 0 2>&1 | tail -1
  done
done
echo "== 296 streams"; for P in "9,9,9,9,9,9" "0,5,4,3,2,1"; do FVAD_PRIO=$P python tools/perf_probe.py 296 120 2 1 0 2>&1 | tail -1; done
echo "== 148 streams"; for P in "9,9,9,9,9,9" "0,5,4,3,2,1"; do FVAD_PRIO=$P python tools/perf_probe.py 148 120 2 1 0 2>&1 | tail -1; done
FVAD_TIMELINE=gpurun_out/timeline_592.csv python tools/perf_probe.py 592 60 2 1 1 2>&1 | head -2
} 2>&1 | tee gpurun_out/prio.log
