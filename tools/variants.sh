#!/bin/bash
# build tuning variants of the library: tools/variants.sh name "-DPK_X=1 ..." [name flags]...
mkdir -p build/variants
S=empower-srslte_b200/csrc
while [ $# -ge 2 ]; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xcompiler -fPIC -shared $2 -o build/variants/$1.so $S/fec_kernels.cu $S/fec_tdec_packed.cu $S/fec_frontend.cu $S/fec_tx.cu $S/fec_cuda_api.cu $S/fec_compat.cu $S/fec_noalloc.cu $S/fec_tables.c 2>&1 | grep -E "error" &
  shift 2
done
wait; ls -la build/variants
