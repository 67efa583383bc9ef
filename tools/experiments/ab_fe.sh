#!/bin/bash
# A/B of the symbol legs (front end in front of the decoder): in-tree library and every build/variants/*.so
run() { FEC_CUDA_LIB=$1 timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-tx --no-harq --sustain 0 --no-worst-case --no-grid 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); f=j['frontend_symbols']; u=j['ulsch_config5']; l=j['llr_8bit'].get('symbols_in') or {}
print('$2', 'DL syms %.2f Gbit/s (step %.3f ms, fe %.3f) e2e %.2f | UL %.2f Gbit/s (step %.3f, fe %.3f) | int8 syms %.2f (fe %.3f)' % (f['value'], f['ms_per_step'], f['serial_phases_ms']['frontend'], f['e2e']['value'], u['value'], u['ms_per_step'], u['serial_phases_ms_rank0']['frontend'], l.get('value', 0), l.get('frontend_ms', 0)))"; }
run "" in-tree
for f in build/variants/*.so; do run $PWD/$f $f; done
run "" in-tree
