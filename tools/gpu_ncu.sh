#!/bin/bash
# One parameterised ncu recipe (B200_PROFILING.md): launch list of a short bench run + one full capture per kernel regex.
#   tools/gpu_ncu.sh [-l] [-b "<bench args>"] NAME=REGEX[:SKIP] ...
#   -l            also write the launch list (gpu__time_duration.sum per launch) to gpurun_out/launches.csv
#   -b "<args>"   extra bench.py arguments (default: the headline leg only)
#   NAME=REGEX:N  full capture (--set full, source counters) of the N-th launch matching REGEX -> gpurun_out/prof_NAME.ncu-rep
# Every ncu pass runs only after the same command has exited 0 without ncu. Numbers printed under ncu are never bench values.
mkdir -p gpurun_out
LIST=0; BARGS="--no-cpu-baseline --no-worst-case --no-symbols --no-tx --no-llr8"
while getopts "lb:" o; do case $o in l) LIST=1;; b) BARGS="$OPTARG";; esac; done
shift $((OPTIND - 1))
CMD="python bench.py --steps 2 --warmup 3 $BARGS"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
[ $LIST = 1 ] && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
for spec in "$@"; do
  name=${spec%%=*}; rest=${spec#*=}; pat=${rest%%:*}; skip=0; [[ $rest == *:* ]] && skip=${rest##*:}
  ncu --set full --clock-control none --import-source on -k regex:"$pat" -s $skip -c 1 -f -o gpurun_out/prof_$name $CMD > gpurun_out/ncu_$name.log 2>&1
  tail -2 gpurun_out/ncu_$name.log
done
ls -la gpurun_out | grep ncu-rep
