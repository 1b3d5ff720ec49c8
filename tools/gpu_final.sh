#!/bin/bash
# final verification of a round: GPU tests, smoke, bench (both arms)
out=gpurun_out/${1:-final}; mkdir -p $out
(time timeout 1300 python -m pytest tests -m gpu -q --durations=5) > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/pytest_gpu.log
tail -12 $out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee $out/smoke.log
python bench.py > $out/bench_n1.json 2> $out/bench_n1.err; echo "bench rc=$?"; cut -c1-600 $out/bench_n1.json
python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_ref.json 2> $out/bench_ref.err; echo "ref rc=$?"; cut -c1-400 $out/bench_ref.json
