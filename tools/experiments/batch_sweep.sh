#!/bin/bash
# batch / sub-batch sweep of the device-resident arm: tools/experiments/batch_sweep.sh [lib.so]
F="--no-cpu-baseline --no-symbols --no-llr8 --no-tx --no-harq --sustain 0 --no-worst-case"
for cfg in ${CFGS:-1184_592 2368_1184 4736_2368}; do
  set -- ${cfg/_/ }
  for lib in "" $LIBS; do
  FEC_CUDA_LIB=$lib FEC_CUDA_SUB_TBS_DEV=$2 python bench.py --batch $1 $F 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 $2', '$lib'[-12:], round(d['value'],2), round(d['ms_per_step'],3), round(d['sync_ms_per_step'],3), d['roofline']['phases_ms'])"
  done
done
