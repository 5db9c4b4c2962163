#!/bin/bash
# one GPU round trip while tuning the fused kernel: parity tests, kernel / step timing, instruction count
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python profiles/kbench.py "$1" 2>&1 | tail -2
python profiles/stepbench.py 2>&1 | tail -5
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum --clock-control none \
  -k regex:fbank400_v3 -s 2 -c 1 python profiles/prof_fbank.py 4 2>&1 | grep -E "inst_executed|time_duration|wavefronts"
ncu --section SourceCounters --section WarpStateStats --import-source on --clock-control none \
  -k regex:fbank400_v3 -s 2 -c 1 -f -o gpurun_out/fbank_q python profiles/prof_fbank.py 4 > /dev/null 2>&1
