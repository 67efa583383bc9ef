#!/bin/bash
# ncu launch list + full capture of one kernel (regex $1, default fec_tdec_pk) of the bench workload at its real size
mkdir -p gpurun_out
PAT=${1:-fec_tdec_pk}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 8 -c 1 -o gpurun_out/prof_map $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/plain.log | cut -c1-300; tail -3 gpurun_out/ncu2.log
