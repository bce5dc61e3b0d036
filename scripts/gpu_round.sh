#!/bin/bash
# One GPU-box visit: parity tests, bench (both arms), ncu launch list, one --set full capture of the top kernel.
# usage (from the repo root, on the GPU box):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/${TAG}_smi.txt 2>&1
nproc > $OUT/${TAG}_nproc.txt; free -g >> $OUT/${TAG}_nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
tail -c 3000 $OUT/${TAG}_bench.json
if [ "${SKIP_NCU:-0}" != "1" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 60 -c 6 -o $OUT/${TAG}_gemm_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
