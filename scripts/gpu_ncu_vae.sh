#!/bin/bash
# ncu --set full of the last-level VAE convolutions (k7 + Snake, then k1 + residual) as tap-GEMM launches
set -x
python scripts/trace_vae.py 30 > gpurun_out/ncu_vae_plain.log 2>&1 || exit 1
ACE_B200_PDL=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel --launch-skip 8 --launch-count 2 \
    -o gpurun_out/r1_vae_conv_last -f python scripts/trace_vae.py 30 > gpurun_out/ncu_vae.log 2>&1
tail -5 gpurun_out/ncu_vae.log
