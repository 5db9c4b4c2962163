#!/bin/bash
# one GPU round trip while tuning the fused kernel: parity tests, kernel / step timing, counters
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python profiles/kbench.py "$1" 2>&1 | tail -2
KBENCH_PADDED=0 python profiles/kbench.py "$1 unpadded" 2>&1 | tail -1
python profiles/stepbench.py 2>&1 | tail -1
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none \
  -k regex:fbank400_v3 -s 2 -c 1 python profiles/prof_fbank.py 4 2>&1 | grep -E "inst_executed|time_duration|wavefronts|bank_conflicts|issue_active"
