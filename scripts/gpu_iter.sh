#!/bin/bash
# Quick iteration on the GPU box: parity tests, then the large-M probes (no ncu).
OUT=gpurun_out; TAG=${1:-it}
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
timeout 900 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
timeout 900 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
