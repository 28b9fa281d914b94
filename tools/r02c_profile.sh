#!/bin/bash
# Round-2 evidence for the final code (CTA-pair convolutions): ncu launch list, one full capture of the paired conv,
# the other BASELINE configs.  Everything lands in gpurun_out/.
set -x
O=gpurun_out
mkdir -p $O
timeout 100 python tools/profile_unet.py --batch 64 > $O/plain_unet.log 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --nvtx --nvtx-include "fwd/" --csv \
    --log-file $O/r02_launches_unet_b64.csv python tools/profile_unet.py --batch 64 > $O/ncu_unet.log 2>&1
timeout 60 python tools/gemm_bench.py 524288 320 320 --taps 9 --hw 64 > $O/plain_conv.log 2>&1 && \
timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 3 -c 1 \
    -o $O/r02_ncu_full_gemm_tc_conv3x3_c320_cg2 python tools/gemm_bench.py 524288 320 320 --taps 9 --hw 64 > $O/ncu_conv.log 2>&1
timeout 300 python tools/config_bench.py > $O/r02_configs_n1.jsonl 2> $O/r02_configs.err
timeout 300 python tools/glide_bench.py --batch 32 > $O/r02_glide_cascade_b32_n1.json 2> $O/r02_glide.err
timeout 120 python tools/gemm_breakdown.py > $O/r02_gemm_tc_breakdown_b64.txt 2>&1
ls -la $O | tail -12
