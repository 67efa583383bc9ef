#!/bin/bash
# worst-case leg only (every block runs 8 half-iterations: the decoder's pure throughput), in-tree and build/variants/*.so
run() { FEC_CUDA_LIB=$1 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-symbols --no-tx --no-llr8 --no-harq --sustain 0 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); w=j['worst_case_8_halfits']; print('$2', 'worst %.3f ms (%.2f Gbit/s, %.1f M CB/s)' % (w['ms_per_step'], w['gbps'], w['cbs_per_s'] / 1e6))"; }
run "" in-tree
for f in build/variants/*.so; do run $PWD/$f $f; done
