#!/bin/bash
# device timeline of the sub-batches of a few steps (FEC_CUDA_TRACE)
mkdir -p gpurun_out
FEC_CUDA_TRACE=1 timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-worst-case > gpurun_out/trace.json 2> gpurun_out/trace.err
grep "fec trace" gpurun_out/trace.err | sed -n '19,42p'
