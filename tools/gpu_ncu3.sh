#!/bin/bash
# ncu launch list + full captures of the current kernels: MAP (a-priori and certificate variants), rate de-matcher, TX encoder
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fec_tdec_pk_kernel<16, true, false>" -s 8 -c 1 -o gpurun_out/prof_map $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fec_tdec_pk_kernel<16, true, true>" -s 4 -c 1 -o gpurun_out/prof_trk $CMD > gpurun_out/ncu5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fec_rm_rx -s 8 -c 1 -o gpurun_out/prof_rm $CMD > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fec_tx_encode -s 2 -c 1 -o gpurun_out/prof_tx $CMD > gpurun_out/ncu4.log 2>&1
tail -3 gpurun_out/ncu2.log; tail -2 gpurun_out/ncu5.log; tail -2 gpurun_out/ncu3.log; tail -2 gpurun_out/ncu4.log; ls -la gpurun_out | grep ncu-rep
