#!/bin/bash
# Large-M probe: GEMM shape sweep + config-3-like sampler (Q4_K, 240 s, batch 2 x CFG) + 240 s VAE, then ncu --set full on the big GEMMs.
OUT=gpurun_out; TAG=${1:-big}
mkdir -p $OUT
timeout 900 python scripts/gemm_sweep.py 0,12 > $OUT/${TAG}_sweep.log 2>&1; echo "sweep exit $?"
timeout 900 python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 8 --vae-seconds 240 --reps 2 > $OUT/${TAG}_probe_q4k.log 2>&1; echo "probe exit $?"
tail -30 $OUT/${TAG}_probe_q4k.log
timeout 900 python scripts/perf_probe.py --quant F16 --seconds 240 --batch 2 --cfg --steps 8 --reps 2 > $OUT/${TAG}_probe_f16.log 2>&1; echo "probe f16 exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 200 -c 8 -o $OUT/${TAG}_gemm_q4k_full -f \
  python scripts/perf_probe.py --quant Q4_K --seconds 240 --batch 2 --cfg --steps 2 --reps 1 > $OUT/${TAG}_ncu_q4k.log 2>&1; echo "ncu exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 200 -c 8 -o $OUT/${TAG}_gemm_f16_full -f \
  python scripts/perf_probe.py --quant F16 --seconds 240 --batch 2 --cfg --steps 2 --reps 1 > $OUT/${TAG}_ncu_f16.log 2>&1; echo "ncu exit $?"
