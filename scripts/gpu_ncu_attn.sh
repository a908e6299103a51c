#!/bin/bash
# ncu --set full of ONE chunk-versus-cache attention launch of the batched step (8 streams x 8192 keys)
export AHA_STREAMS=${AHA_STREAMS:-8} AHA_WINDOW=${AHA_WINDOW:-8192} AHA_CACHE=${AHA_CACHE:-sliding_window}
TAG=${1:-attn}
python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_plain.log 2>&1 &&
AHA_NCU=1 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attention_persistent -s 3 -c 1 \
    -o gpurun_out/${TAG} python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_ncu.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/${TAG}_plain.log; ls -la gpurun_out/${TAG}.ncu-rep
