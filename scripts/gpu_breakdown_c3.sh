#!/bin/bash
# in-situ cost of each per-layer kernel of a batched config-3 decoder step: step time with one kernel skipped (AHA_SKIP bit)
export AHA_STREAMS=${AHA_STREAMS:-8} AHA_WINDOW=8192 AHA_CACHE=sliding_window
for pair in ${PAIRS:-1 0}; do
  for skip in ${SKIPS:-0 8 16 32 64 4 2 126}; do
    AHA_PS_PAIR=$pair AHA_SKIP=$skip python scripts/gpu_decoder_breakdown.py 32 2>&1 | tail -1
  done
done
