#!/bin/bash
OUT=gpurun_out; TAG=${1:-pair}
mkdir -p $OUT
timeout 180 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "cta_pair" > $OUT/${TAG}_pytest_pair.log 2>&1; rc=$?; echo "pair pytest exit $rc"; tail -15 $OUT/${TAG}_pytest_pair.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
timeout 600 python scripts/gemm_sweep.py 0 > $OUT/${TAG}_sweep.log 2>&1; echo "sweep exit $?"; cut -c1-330 $OUT/${TAG}_sweep.log
timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 2 > $OUT/${TAG}_probe_f16_60.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_f16_60.log
