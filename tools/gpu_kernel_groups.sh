#!/bin/bash
# run kernel test groups in separate processes (a trapped kernel poisons its CUDA context)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
for grp in "linear" "conv3x3" "geglu or qkv" "attention" "groupnorm or layernorm or softmax" "cfg or timestep or resample or small or u8"; do
  echo "=== $grp" 
  timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "$grp" -p no:cacheprovider 2>&1 | tail -25
done
