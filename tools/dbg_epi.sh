#!/bin/bash
# development: conv-class time with parts of the transposing epilogue switched off (results are garbage, timing only)
for d in 0 1 2 3; do echo "== DMX_DBG_EPI=$d"; DMX_DBG_EPI=$d python tools/profile_step.py 2>&1 | grep -E "gemm_conv|device ms"; done
