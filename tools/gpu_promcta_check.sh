#!/bin/bash
out=gpurun_out/${1:-promcta}; mkdir -p $out
(timeout 900 python -m pytest tests/test_gpu_prom.py tests/test_gpu_determinism.py -m gpu -q -x 2>&1 | tail -5) > $out/pytest.txt
tail -3 $out/pytest.txt
for a in "4096 40 LSPG 100" "4096 40 Galerkin 100" "4096 47 Galerkin 100" "2025 64 LSPG 100" "2025 64 Galerkin 100" "1024 96 LSPG 100"; do
  timeout 300 python tools/prom_time.py $a 2>&1 | tail -1
done | tee $out/time.txt
for t in "8,8" "8,12" "8,16" "8,32"; do echo "tune=$t"; for a in "4096 40 Galerkin 50"; do BX_PROM_CTA_TUNE=$t timeout 300 python tools/prom_time.py $a 2>&1 | tail -1; done; done | tee $out/tune.txt
