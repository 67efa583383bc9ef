#!/bin/bash
# launch list of a whole bench run (all legs) + full captures of the MAP kernel and of the TB assembly kernel
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"pk_kernel<.int.16, .bool.1, .bool.0>" -s 8 -c 1 -o gpurun_out/prof_map2 $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --kernel-name-base demangled -k regex:"fec_tb_assemble" -s 8 -c 1 -o gpurun_out/prof_asm $CMD > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"win_kernel<.int.32, .bool.1>" -s 4 -c 1 -o gpurun_out/prof_win8b $CMD > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu2.log | cut -c1-200; ls -la gpurun_out | grep ncu-rep
