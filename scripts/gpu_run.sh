#!/bin/bash
# One parameterised GPU-box script (replaces round 1's sixteen gpu_*.sh): every stage writes under gpurun_out/<tag>_*.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_run.sh <tag> <stage> [<stage> ...]'
# stages: tests[:<pytest -k expr>]  smoke  bench[:<extra bench.py args>]  bench_ref  launches[:<bench args>]  ncu:<kernel regex>[:<bench args>]
#         sh:<shell command>   torchrun:<N>[:<bench args>]   probe:<perf_probe.py args>   sass
OUT=gpurun_out; TAG=${1:-run}; shift
mkdir -p $OUT
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1
for st in "$@"; do
  kind=${st%%:*}; arg=""; [[ "$st" == *:* ]] && arg=${st#*:}
  t0=$(date +%s)
  case $kind in
    tests)   timeout 1500 python -m pytest tests -x -q -m gpu ${arg:+-k "$arg"} --durations=15 > $OUT/${TAG}_pytest.log 2>&1; rc=$?; tail -25 $OUT/${TAG}_pytest.log ;;
    smoke)   timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; rc=$?; tail -3 $OUT/${TAG}_smoke.log ;;
    bench)   name=$(echo "b${arg}" | tr -c 'A-Za-z0-9\n' '_'); timeout 1200 python bench.py $arg > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.log; rc=$?; cat $OUT/${TAG}_${name}.json; tail -5 $OUT/${TAG}_${name}.log ;;
    torchrun) n=${arg%%:*}; bargs=""; [[ "$arg" == *:* ]] && bargs=${arg#*:}; name=$(echo "n${n}${bargs}" | tr -c 'A-Za-z0-9\n' '_')
             timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n $bargs \
                 > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.log; rc=$?; cat $OUT/${TAG}_${name}.json; tail -4 $OUT/${TAG}_${name}.log ;;
    bench_ref) timeout 900 python bench.py --impl reference --steps 1 $arg > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.log; rc=$?; cat $OUT/${TAG}_bench_ref.json ;;
    launches) timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/${TAG}_launches.csv \
                 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song $arg > $OUT/${TAG}_launches.log 2>&1; rc=$? ;;
    ncu)     kre=${arg%%:*}; bargs=""; [[ "$arg" == *:* ]] && bargs=${arg#*:}
             timeout 1500 ncu --set full --clock-control none -k regex:$kre -s 40 -c 3 -o $OUT/${TAG}_ncu_$(echo $kre | tr -c 'A-Za-z0-9\n' '_') -f \
                 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-song $bargs > $OUT/${TAG}_ncu.log 2>&1; rc=$? ;;
    probe)   timeout 900 python scripts/perf_probe.py $arg > $OUT/${TAG}_probe.log 2>&1; rc=$?; tail -20 $OUT/${TAG}_probe.log ;;
    sh)      name=$(echo "sh_${arg}" | tr -c 'A-Za-z0-9\n' '_' | cut -c1-60); timeout 1200 bash -c "$arg" > $OUT/${TAG}_${name}.log 2>&1; rc=$?; tail -12 $OUT/${TAG}_${name}.log ;;
    sass)    bash scripts/sass_summary.sh > $OUT/${TAG}_sass.txt 2>&1; rc=$? ;;
    *)       echo "unknown stage $st"; rc=99 ;;
  esac
  echo "== stage $st exit $rc ($(( $(date +%s) - t0 )) s)"
done
