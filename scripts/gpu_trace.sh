#!/bin/bash
OUT=gpurun_out; TAG=${1:-tr}
mkdir -p $OUT
cat > /tmp/trace_one.py <<'PY'
import os, sys
import numpy as np
sys.path[:0] = ["/root/repo", "/root/repo/ace-step-1.5-ggml_b200", "/root/repo/tests"]
from acestep_b200.bridge import B200Bridge
br = B200Bridge()
g = np.random.default_rng(0)
os.environ["ACE_B200_GEMM_2CTA"] = "0"
for (M, N, K, bn, sk) in [(750, 2048, 2048, 128, 1), (750, 2048, 2048, 128, 3), (750, 2048, 2048, 256, 2), (750, 4096, 2048, 256, 1), (750, 12288, 2048, 256, 1), (750, 2048, 6144, 256, 3), (6000, 2048, 2048, 256, 1)]:
    A = g.standard_normal((1, M, K), dtype=np.float32)
    W = g.standard_normal((N, K), dtype=np.float32) * 0.02
    R = np.zeros((1, M, N), dtype=np.float32)
    os.environ["ACE_B200_GEMM_BN"] = str(bn); os.environ["ACE_B200_GEMM_SPLITK"] = str(sk)
    os.environ.pop("ACE_B200_GEMM_TRACE", None)
    for _ in range(3):
        br.debug_gemm(A, W, w_type=0, N=N, resid=R)
    print(f"M={M} N={N} K={K} bn={bn} sk={sk}: {br.last_device_ms()*1e3:.1f} us", flush=True)
    os.environ["ACE_B200_GEMM_TRACE"] = "1"
    br.debug_gemm(A, W, w_type=0, N=N, resid=R)
br.close()
PY
timeout 300 python /tmp/trace_one.py > $OUT/${TAG}_trace.log 2>&1; echo "trace exit $?"; cat $OUT/${TAG}_trace.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_tc_kernel -s 30 -c 4 -o $OUT/${TAG}_attn_full -f \
  python scripts/perf_probe.py --quant F16 --seconds 240 --batch 2 --cfg --steps 1 --reps 1 > $OUT/${TAG}_ncu_attn.log 2>&1; echo "ncu attn exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_vae_launches.csv \
  python scripts/perf_probe.py --seconds 0 --vae-seconds 240 --reps 1 > $OUT/${TAG}_ncu_vae.log 2>&1; echo "ncu vae exit $?"
