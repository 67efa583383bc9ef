#!/bin/bash
# sweep of the sub-batch size (device-resident arm)
for sub in ${SWEEP:-400 500 592 700 800}; do
  FEC_CUDA_SUB_TBS_DEV=$sub timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-worst-case 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('sub $sub: pipelined %.3f sync %.3f e2e %.3f / %.3f' % (j['ms_per_step'], j['sync_ms_per_step'], j['e2e']['ms_per_step'], j['e2e']['sync_ms_per_step']))"
done
