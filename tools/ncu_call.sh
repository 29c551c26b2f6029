#!/bin/bash
# usage: ncu_call.sh <targets> ; run on the GPU box from the repo root
W=$1
TARGETS=$W python benchmarks/ncu_targets.py > gpurun_out/ncu_plain_run_$W.log 2>&1; rc=$?; echo "targets $W exit $rc"; tail -2 gpurun_out/ncu_plain_run_$W.log
if [ $rc -eq 0 ]; then
  TARGETS=$W timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/prof_r2_$W python benchmarks/ncu_targets.py > gpurun_out/ncu_run_$W.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_run_$W.log
  ncu -i gpurun_out/prof_r2_$W.ncu-rep --page raw --csv > gpurun_out/prof_r2_${W}_raw.csv 2>/dev/null
  ls -la gpurun_out/
  sz=$(stat -c %s gpurun_out/prof_r2_$W.ncu-rep); if [ $sz -gt 50000000 ]; then echo "report too large to bring back ($sz bytes): keeping the csv pages only"; ncu -i gpurun_out/prof_r2_$W.ncu-rep --page source --csv > gpurun_out/prof_r2_${W}_source.csv 2>/dev/null; gzip -9 gpurun_out/prof_r2_${W}_source.csv; rm gpurun_out/prof_r2_$W.ncu-rep; fi
fi
