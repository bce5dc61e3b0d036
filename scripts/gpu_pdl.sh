#!/bin/bash
OUT=gpurun_out; TAG=${1:-pdl}
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -6 $OUT/${TAG}_pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
for pdl in 1 0; do
ACE_B200_PDL=$pdl timeout 600 python scripts/perf_probe.py --quant F16 --seconds 60 --batch 1 --steps 8 --reps 3 > $OUT/${TAG}_probe_f16_60_pdl$pdl.log 2>&1; echo "probe pdl=$pdl exit $?"
grep -E "^\[(dit|abi)" $OUT/${TAG}_probe_f16_60_pdl$pdl.log
done
ACE_B200_PDL=1 timeout 600 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
grep -E "^\[(dit|prof|vae|abi)" $OUT/${TAG}_probe_q4k.log
