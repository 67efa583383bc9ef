#!/bin/bash
mkdir -p gpurun_out
for v in 296 592 1184; do
FEC_CUDA_SUB_TBS_DEV=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-worst-case 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('dev',$v, j['value'], j['ms_per_step'], j['e2e']['value'], j['e2e']['ms_per_step'])"
done
for v in 148 296 592; do
FEC_CUDA_SUB_TBS=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-worst-case 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('host',$v, j['value'], j['ms_per_step'], j['e2e']['value'], j['e2e']['ms_per_step'])"
done
