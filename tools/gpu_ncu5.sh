#!/bin/bash
# full capture of the int8 windowed decoder (8-bit LLR leg of the bench) and of the int8 rate de-matcher at full size
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-worst-case --no-tx --no-symbols"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"win_kernel<.int.32, .bool.1>" -s 4 -c 1 -o gpurun_out/prof_win8 $CMD > gpurun_out/ncu8.log 2>&1
ncu --set full --clock-control none --kernel-name-base demangled -k regex:"rm_rx8_kernel" -s 4 -c 1 -o gpurun_out/prof_rm8 $CMD > gpurun_out/ncu9.log 2>&1
tail -3 gpurun_out/ncu8.log; tail -3 gpurun_out/ncu9.log; ls -la gpurun_out | grep ncu-rep
