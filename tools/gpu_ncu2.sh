#!/bin/bash
# ncu launch list + full captures (MAP kernel at its real size, rate de-matcher, DL front end) of the bench workload
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fec_tdec_pk -s 8 -c 1 -o gpurun_out/prof_map $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fec_rm_rx -s 8 -c 1 -o gpurun_out/prof_rm $CMD > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fec_fe_dl -s 2 -c 1 -o gpurun_out/prof_fe $CMD > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/plain.log | cut -c1-300; tail -3 gpurun_out/ncu2.log; tail -2 gpurun_out/ncu3.log; tail -2 gpurun_out/ncu4.log; ls -la gpurun_out
