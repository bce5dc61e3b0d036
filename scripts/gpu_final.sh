#!/bin/bash
# Round-end style visit: smoke, bench (both arms), ncu launch list + --set full captures of the top kernels on the bench workload.
OUT=gpurun_out; TAG=${1:-fin}
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/${TAG}_smi.txt 2>&1; nproc >> $OUT/${TAG}_smi.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/${TAG}_smoke.log
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"; cat $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "bench ref exit $?"; cat $OUT/${TAG}_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1300 -c 900 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 800 -c 14 -o $OUT/${TAG}_gemm_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song > $OUT/${TAG}_ncu_gemm.log 2>&1; echo "ncu gemm exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_tcp -s 200 -c 4 -o $OUT/${TAG}_attn_full -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song > $OUT/${TAG}_ncu_attn.log 2>&1; echo "ncu attn exit $?"
timeout 300 python scripts/trace_vae.py 240 2>&1 | head -2 > $OUT/${TAG}_vae240.log; cat $OUT/${TAG}_vae240.log
