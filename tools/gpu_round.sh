#!/bin/bash
# One GPU session: GPU test tier, smoke, our bench, the reference bench.  Logs -> gpurun_out/round/
cd "$(dirname "$0")/.."
OUT=gpurun_out/round
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $OUT/summary.txt
timeout 600 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/summary.txt
timeout 900 python bench.py --steps ${STEPS:-20} --warmup 5 > $OUT/bench_ours.json 2> $OUT/bench_ours.err; echo "bench ours rc=$?" | tee -a $OUT/summary.txt
cat $OUT/bench_ours.json
if [ "${SKIP_REF:-0}" != "1" ]; then
timeout 1500 python bench.py --impl reference --steps ${STEPS:-20} --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?" | tee -a $OUT/summary.txt
cat $OUT/bench_ref.json
fi
tail -5 $OUT/pytest_gpu.log
