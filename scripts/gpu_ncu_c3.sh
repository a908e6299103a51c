#!/bin/bash
# ncu launch list of ONE steady-state batched decoder step (config 3 shapes), after the same command ran clean without ncu
export AHA_STREAMS=${AHA_STREAMS:-8} AHA_WINDOW=8192 AHA_CACHE=sliding_window AHA_PS_PAIR=${AHA_PS_PAIR:-0}
TAG=${1:-c3}
python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_plain.log 2>&1 &&
AHA_NCU=1 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/${TAG}_launches.csv python scripts/gpu_decoder_breakdown.py 8 > gpurun_out/${TAG}_ncu.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/${TAG}_plain.log
python scripts/summarize_launches.py gpurun_out/${TAG}_launches.csv 1 20 36 2>&1 | tail -40
