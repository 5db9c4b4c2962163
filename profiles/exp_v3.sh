#!/bin/bash
# builds experiment variants of the fused kernel ON the GPU box and times each
for v in "" "-DSBFE_EXP_SKIP4" "-DSBFE_EXP_SKIP34"; do
  SBFE_NVCC_EXTRA="$v" python -m speechbrain_b200._build --force > /dev/null 2>&1
  python profiles/kbench.py "[$v]" 2>&1 | tail -2
done
python -m speechbrain_b200._build --force > /dev/null 2>&1
