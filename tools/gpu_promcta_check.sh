#!/bin/bash
mkdir -p gpurun_out/r02x
(timeout 120 python -m pytest "tests/test_gpu_prom.py::test_large_basis_vs_oracle" -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r02x/pytest0.txt
tail -15 gpurun_out/r02x/pytest0.txt
grep -q "failed\|error\|Error" gpurun_out/r02x/pytest0.txt && exit 1
grep -q "passed" gpurun_out/r02x/pytest0.txt || exit 1
(timeout 600 python -m pytest tests/test_gpu_prom.py tests/test_gpu_determinism.py -m gpu -q -x 2>&1 | tail -5) > gpurun_out/r02x/pytest.txt
tail -3 gpurun_out/r02x/pytest.txt
for a in "4096 40 LSPG 100" "4096 40 Galerkin 100" "2025 64 LSPG 100" "2025 64 Galerkin 100" "1024 96 LSPG 100" "1024 96 Galerkin 100" "289 227 LSPG 50" "289 227 Galerkin 50"; do
  timeout 300 python tools/prom_time.py $a 2>&1 | tail -1
done | tee gpurun_out/r02x/time.txt
for t in "8,28" "8,20" "16,28"; do echo "tune=$t"; for a in "4096 40 LSPG 50" "2025 64 LSPG 50" "1024 96 Galerkin 50"; do BX_PROM_CTA_TUNE=$t timeout 300 python tools/prom_time.py $a 2>&1 | tail -1; done; done | tee gpurun_out/r02x/tune.txt
