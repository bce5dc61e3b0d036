#!/bin/bash
# ncu --set full of the QKV projection (EPI_HEADNORM) and the o-projection (TMA reduce-add) on the bench workload
ACE_B200_PDL=0 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:gemm_tc_kernel<\(int\)256' --launch-skip 60 --launch-count 2 -o gpurun_out/r1_dit_qkv -f python scripts/trace_dit.py > gpurun_out/ncu_hn.log 2>&1
tail -3 gpurun_out/ncu_hn.log
