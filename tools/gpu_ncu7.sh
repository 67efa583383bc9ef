#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case --no-tx --no-symbols"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"win8p_kernel<.int.32, .bool.1>" -s 4 -c 1 -o gpurun_out/prof_win8p $CMD > gpurun_out/ncu8.log 2>&1
tail -2 gpurun_out/ncu8.log | cut -c1-200; ls -la gpurun_out | grep win8p
