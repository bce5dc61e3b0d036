#!/bin/bash
OUT=gpurun_out; TAG=${1:-tr}
mkdir -p $OUT
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "gemm" > $OUT/${TAG}_pytest_gemm.log 2>&1; rc=$?; echo "gemm pytest exit $rc"; tail -12 $OUT/${TAG}_pytest_gemm.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -5 $OUT/${TAG}_pytest.log
timeout 300 python scripts/trace_one.py > $OUT/${TAG}_trace.log 2>&1; echo "trace exit $?"; grep -v non-atomic $OUT/${TAG}_trace.log | cut -c1-330 | head -16
for t in 1 0; do
ACE_B200_GEMM_TMARED=$t timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 3 > $OUT/${TAG}_probe_f16_60_t$t.log 2>&1; echo "probe tmared=$t exit $?"
grep -E "^\[(dit|abi)" $OUT/${TAG}_probe_f16_60_t$t.log
done
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
