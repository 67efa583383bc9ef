#!/bin/bash
# A/B of the hand-over variants on one box: serial decode phase and pipelined step of the headline workload, uplink leg
for v in v_cur v_nocont v_norefill v_inl v_inl_norefill v_cur v_nocont; do
  FEC_CUDA_LIB=build/variants/$v.so python bench.py --steps 20 --warmup 5 --no-tx --no-llr8 --no-worst-case --no-cpu-baseline > gpurun_out/ab_$v.json 2>/dev/null
  echo "== $v"; python tools/show_bench.py gpurun_out/ab_$v.json | grep -E "^value|^uplink" | cut -c1-230
done
