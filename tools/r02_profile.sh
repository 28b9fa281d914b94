#!/bin/bash
# Round-2 evidence run on one B200 (everything lands in gpurun_out/; summaries are copied to profiles/ by hand).
set -x
O=gpurun_out
timeout 400 python bench.py --steps 5 --warmup 3 > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; echo "bench rc=$?"
timeout 120 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench_reference.json 2>> $O/r02_bench_n1.err
timeout 300 python tools/config_bench.py > $O/r02_configs_n1.jsonl 2> $O/r02_configs.err
timeout 300 python tools/glide_bench.py --batch 32 > $O/r02_glide_cascade_b32_n1.json 2> $O/r02_glide.err
timeout 120 python tools/gemm_breakdown.py > $O/r02_gemm_tc_breakdown_b64.txt 2>&1
timeout 120 python tools/norm_bench.py r02 > $O/r02_norm_bench.txt 2>&1
timeout 60 python tools/attn_bench.py 32 5 4096 4096 64 > $O/r02_attn_bench.txt 2>&1
timeout 100 python tools/profile_unet.py --batch 64 > $O/plain_unet.log 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --nvtx --nvtx-include "fwd/" --csv \
    --log-file $O/r02_launches_unet_b64.csv python tools/profile_unet.py --batch 64 > $O/ncu_unet.log 2>&1
timeout 60 python tools/gemm_bench.py 524288 320 320 --taps 9 --hw 64 > $O/plain_conv.log 2>&1 && \
timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 3 -c 1 \
    -o $O/r02_ncu_full_gemm_tc_conv3x3_c320 python tools/gemm_bench.py 524288 320 320 --taps 9 --hw 64 > $O/ncu_conv.log 2>&1
timeout 60 python tools/gemm_prof.py to_out > $O/plain_toout.log 2>&1 && \
timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1 -c 1 \
    -o $O/r02_ncu_full_gemm_tc_to_out python tools/gemm_prof.py to_out > $O/ncu_toout.log 2>&1
ls -la $O | tail -30
