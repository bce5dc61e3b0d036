#!/bin/bash
OUT=gpurun_out; TAG=${1:-pair}
mkdir -p $OUT
timeout 180 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "cta_pair" > $OUT/${TAG}_pytest_pair.log 2>&1; rc=$?; echo "pair pytest exit $rc"; tail -5 $OUT/${TAG}_pytest_pair.log
if [ $rc -ne 0 ]; then exit 0; fi
ACE_B200_DEBUG=1 timeout 600 python scripts/gemm_sweep.py 0 > $OUT/${TAG}_sweep.log 2>&1; echo "sweep exit $?"; cut -c1-130,250-400 $OUT/${TAG}_sweep.log
