#!/bin/bash
# A/B on one box: the in-tree library and every build/variants/*.so, serial kernel phases + pipelined step + worst case
mkdir -p gpurun_out
run() { FEC_CUDA_LIB=$1 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-symbols --no-tx --no-llr8 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); r=j['roofline']['phases_ms']; w=j['worst_case_8_halfits']
print('$2', 'dec %.3f rm %.3f asm %.3f step %.3f (%.1f Gbit/s) worst %.3f (%.1f)' % (r['turbo_decode'], r['rate_dematch'], r['tb_assemble_crc24a'], j['ms_per_step'], j['value'], w['ms_per_step'], w['gbps']))"; }
run "" in-tree
for f in build/variants/*.so; do run $PWD/$f $f; done
run "" in-tree
