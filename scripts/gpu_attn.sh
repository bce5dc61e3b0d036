#!/bin/bash
OUT=gpurun_out; TAG=${1:-attn}
mkdir -p $OUT
timeout 240 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $OUT/${TAG}_pytest_attn.log 2>&1; rc=$?; echo "attn pytest exit $rc"; tail -25 $OUT/${TAG}_pytest_attn.log
if [ $rc -ne 0 ]; then
  timeout 240 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "attention" > $OUT/${TAG}_pytest_attn_all.log 2>&1; grep -E "PASS|FAIL|passed|failed|rel_err|assert" $OUT/${TAG}_pytest_attn_all.log | head -40
  exit 0
fi
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
