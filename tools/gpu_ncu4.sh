#!/bin/bash
# full captures of the two packed MAP kernel variants at full size (demangled names carry the template arguments)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case --no-tx"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"pk_kernel<.int.16, .bool.1, .bool.0>" -s 8 -c 1 -o gpurun_out/prof_map $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"pk_kernel<.int.16, .bool.1, .bool.1>" -s 6 -c 1 -o gpurun_out/prof_trk $CMD > gpurun_out/ncu5.log 2>&1
tail -3 gpurun_out/ncu2.log; tail -3 gpurun_out/ncu5.log; ls -la gpurun_out | grep ncu-rep
