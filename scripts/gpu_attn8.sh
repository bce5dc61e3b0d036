#!/bin/bash
OUT=gpurun_out; TAG=${1:-a8}
mkdir -p $OUT
timeout 240 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $OUT/${TAG}_pytest_attn.log 2>&1; rc=$?; echo "attn pytest exit $rc"; tail -8 $OUT/${TAG}_pytest_attn.log
timeout 300 python scripts/trace_one.py > $OUT/${TAG}_trace.log 2>&1; echo "trace exit $?"; cut -c1-330 $OUT/${TAG}_trace.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
ACE_B200_ATTN=tc4 timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --reps 1 > $OUT/${TAG}_probe_q4k_tc4.log 2>&1; echo "probe tc4 exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k_tc4.log
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
