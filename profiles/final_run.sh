#!/bin/bash
# Round-end measurement pass on one B200: bench (both arms), launch list, ncu --set full of the
# step's kernels, per-config timings, stage ablation.  Everything lands in gpurun_out/.
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2 > $O/final_tests.txt
python bench.py > $O/final_bench.json 2> $O/final_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/final_bench_reference.json 2>> $O/final_bench.err
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/final_bench_short.json 2>> $O/final_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/final_launches.csv \
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/final_ncu_launch.log 2>&1
python profiles/prof_fbank.py 4 > $O/final_prof.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:"fbank400_v3|norm_|clamp_" -s 6 -c 3 -f -o $O/final_step \
  python profiles/prof_fbank.py 4 >> $O/final_prof.log 2>&1
python profiles/stepbench.py > $O/final_stepbench.txt 2>&1
python profiles/configs_bench.py --json $O/final_configs.json > $O/final_configs.txt 2>&1
(python profiles/istft_bench.py; python profiles/nfft512_bench.py) > $O/final_general.txt 2>&1
python profiles/prof_cfg.py cfg3 > /dev/null 2>&1 && ncu --set full --clock-control none -k regex:"post_fast" -s 2 -c 1 -f -o $O/final_post python profiles/prof_cfg.py cfg3 > /dev/null 2>&1
bash profiles/exp_v3.sh > $O/final_ablation.txt 2>&1
python -m speechbrain_b200._build --force > /dev/null 2>&1
cat $O/final_tests.txt $O/final_stepbench.txt $O/final_configs.txt $O/final_ablation.txt
