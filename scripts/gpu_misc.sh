#!/bin/bash
OUT=gpurun_out; TAG=${1:-misc}
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -4 $OUT/${TAG}_pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python scripts/trace_one.py > $OUT/${TAG}_trace.log 2>&1; echo "trace exit $?"; grep -A1 -E "sk=2|SwiGLU" $OUT/${TAG}_trace.log | cut -c1-330
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 3 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|abi)" $OUT/${TAG}_probe_f16_60.log
timeout 600 python scripts/perf_probe.py --seconds 0 --vae-seconds 600 --reps 2 > $OUT/${TAG}_probe_vae600.log 2>&1; echo "vae600 exit $?"
grep -E "^\[(vae|prof)" $OUT/${TAG}_probe_vae600.log
