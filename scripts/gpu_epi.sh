#!/bin/bash
OUT=gpurun_out; TAG=${1:-epi}
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -12 $OUT/${TAG}_pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python /root/repo/scripts/trace_one.py > $OUT/${TAG}_trace.log 2>&1; echo "trace exit $?"; cut -c1-300 $OUT/${TAG}_trace.log
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
