#!/bin/bash
# time every build/variants/*.so with the serial (un-pipelined) kernel phases
for f in build/variants/*.so; do
  FEC_CUDA_LIB=$PWD/$f timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); r=j['roofline']['phases_ms']; w=j['worst_case_8_halfits']
print('$f', 'dec %.3f rm %.3f step %.3f worst %.3f' % (r['turbo_decode'], r['rate_dematch'], j['ms_per_step'], w['ms_per_step']))"
done
