#!/bin/bash
OUT=gpurun_out; TAG=${1:-a3}
mkdir -p $OUT
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "attention" > $OUT/${TAG}_pytest_attn.log 2>&1; rc=$?; echo "attn pytest exit $rc"; grep -E "passed|failed|^FAILED|rel_err" $OUT/${TAG}_pytest_attn.log | head -30
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
if [ $rc -eq 0 ]; then
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
fi
