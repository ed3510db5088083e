#!/bin/bash
# First GPU call of round 2 (one B200, ~60 s of box time): times the candidates left off by default at the end of
# round 1 and checks their bits, then re-times every kernel with the on-stream gs_profile_outer.
#   /usr/local/graft/bin/gpurun --timeout 150 -- 'bash tools/round2_first_call.sh'
# Two GPUs (the staged update inside the peer-memory kernels; charged twice):
#   /usr/local/graft/bin/gpurun --gpus 2 --timeout 300 -- \
#     'GS_TEST_CANDIDATES=1 python -m pytest tests/test_multi_gpu.py -x -q -k "2-peer" > gpurun_out/mgpu_staged.log 2>&1; tail -5 gpurun_out/mgpu_staged.log'
set -u
mkdir -p gpurun_out
timeout 60 python tools/variant_probe.py 0 320 default,asm_nodup C3 > gpurun_out/r2_probe_c3.log 2>&1
timeout 60 python tools/variant_probe.py 0 320 default,asm_nodup C4 > gpurun_out/r2_probe_c4.log 2>&1
grep -h '"us"\|check' gpurun_out/r2_probe_c3.log gpurun_out/r2_probe_c4.log | cut -c1-400
