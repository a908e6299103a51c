#!/bin/bash
# ncu --set full of the four batched decoder GEMM launches of one layer (config-3 shapes): DRAM traffic per launch
export AHA_STREAMS=${AHA_STREAMS:-8} AHA_WINDOW=8192 AHA_CACHE=sliding_window
TAG=${1:-gemm}
python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_plain.log 2>&1 &&
AHA_NCU=1 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_swap_persistent -s 8 -c 4 \
    -o gpurun_out/${TAG} python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_ncu.log 2>&1
echo "rc=$?"; ls -la gpurun_out/${TAG}.ncu-rep
