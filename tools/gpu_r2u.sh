#!/bin/bash
# A/B of the register-kernel variants (reduced solve in registers / shared memory, Phi prefetch depth 3 / 1)
TAG=${1:-r2u}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in "" tools/bin/lib_pw_reg_pf1.so tools/bin/lib_pw_smem_pf3.so tools/bin/lib_pw_smem_pf1.so; do
  echo "variant: ${v:-default (reg LU, pf 3)}"
  for n in 40 32 36 44; do for proj in LSPG Galerkin; do BX_VARIANT_LIB=$v python tools/prom_time.py 4096 $n $proj 60 | tail -1; done; done
done > $OUT/variants.txt 2>&1
cat $OUT/variants.txt
