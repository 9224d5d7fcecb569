#!/bin/bash
# ncu --set full captures of the final build's dominant kernels over tools/profile_step.py (one batch of 16 segments, bf16),
# tabulated on the box with tools/ncu_full_table.py; only the tables (gpurun_out/full_*.md) travel back.
# Run after the plain command exited 0:  python tools/profile_step.py && bash tools/capture_full.sh
set -u
NCU="ncu --profile-from-start off --set full --clock-control none"
run() { # tag, extra ncu args
    tag=$1; shift
    $NCU "$@" -f -o /tmp/full_$tag python tools/profile_step.py > gpurun_out/full_$tag.log 2>&1
    python tools/ncu_full_table.py /tmp/full_$tag.ncu-rep > gpurun_out/full_$tag.md 2>> gpurun_out/full_$tag.log
    rm -f /tmp/full_$tag.ncu-rep
}
run ring_enc_tr  -k regex:gemm_tc_tma -c 14
run ring_dec     -k regex:gemm_tc_tma --launch-skip 51 -c 8
run attn_dconv   -k "regex:attention_tc|istft_frames|stft_kernel|dconv_mma" -c 16
run gather       -k "regex:gemm_tc_kernel" -c 40
wc -l gpurun_out/full_*.md
