#!/bin/bash
mkdir -p gpurun_out
SWEEP_VARIANTS=0,3,8,9,5 python tools/estep_sweep.py 64 2048 4096 2>&1 | grep '^{' | tee gpurun_out/sweep2.jsonl
for IS in 512 2048 4096; do SDMMB_ITEM_SAMPLES=$IS SWEEP_VARIANTS=0 python tools/estep_sweep.py 64 2048 4096 2>&1 | grep '^{' | sed "s/^{/{\"item_samples\": $IS, /" | tee -a gpurun_out/sweep2.jsonl; done
