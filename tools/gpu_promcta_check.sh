#!/bin/bash
mkdir -p gpurun_out/r02u
(timeout 900 python -m pytest tests/test_gpu_prom.py tests/test_gpu_determinism.py -m gpu -q -x 2>&1 | tail -5) > gpurun_out/r02u/pytest.txt
tail -3 gpurun_out/r02u/pytest.txt
for a in "4096 40 LSPG 100" "4096 40 Galerkin 100" "2025 64 LSPG 100" "1024 96 LSPG 100" "289 227 LSPG 50"; do
  timeout 300 python tools/prom_time.py $a 2>&1 | tail -1
done | tee gpurun_out/r02u/time.txt
for t in "8,16" "8,32" "8,8" "16,32"; do echo "tune=$t"; BX_PROM_CTA_TUNE=$t timeout 300 python tools/prom_time.py 4096 40 LSPG 50 2>&1 | tail -1; BX_PROM_CTA_TUNE=$t timeout 300 python tools/prom_time.py 2025 64 LSPG 50 2>&1 | tail -1;  done | tee gpurun_out/r02u/tune.txt
