#!/bin/bash
# Round-2 measurement pass on one B200: tests, bench (both arms), launch list, ncu --set full of the
# step's kernels / the 512-point kernel / the tensor-core projection, per-config timings, timeline.
# Everything lands in gpurun_out/r2f_*; profiles/make_r2_final_md.py assembles profiles/r2_final.md.
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -2 > $O/r2f_tests.txt
python bench.py > $O/r2f_bench.json 2> $O/r2f_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r2f_bench_reference.json 2>> $O/r2f_bench.err
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/r2f_bench_short.json 2>> $O/r2f_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2f_launches.csv \
  python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/r2f_ncu_launch.log 2>&1
python profiles/prof_norm.py > $O/r2f_prof.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:"fbank400_v3|norm_sentence|clamp_tiles" -s 9 -c 3 -f -o $O/r2f_step \
  python profiles/prof_norm.py >> $O/r2f_prof.log 2>&1
PROF_PADDED=0 ncu --set full --import-source on --clock-control none -k regex:fbank400_v3 -s 2 -c 1 -f -o $O/r2f_fbank_unpadded \
  python profiles/prof_fbank.py 4 >> $O/r2f_prof.log 2>&1
ncu --set full --clock-control none -k regex:"fbank512|mel_umma|mel_mma" -s 3 -c 3 -f -o $O/r2f_other python profiles/prof_other.py >> $O/r2f_prof.log 2>&1
python profiles/stepbench.py > $O/r2f_stepbench.txt 2>&1
(for v in 3 4 0; do echo "variant $v (3: FFT kernel alone, 4: + clamp launch, 0: the whole step)"; VARIANT=$v python profiles/twostream.py; done) >> $O/r2f_stepbench.txt 2>&1
python profiles/configs_bench.py --json $O/r2f_configs.json > $O/r2f_configs.txt 2>&1
(python profiles/istft_bench.py; python profiles/stft_bench.py; python profiles/nfft512_bench.py; python profiles/gaussian_bench.py; python profiles/tdbench.py) > $O/r2f_general.txt 2>&1
bash profiles/exp_timeline.sh > $O/r2f_timeline.txt 2>&1
cat $O/r2f_tests.txt $O/r2f_stepbench.txt $O/r2f_configs.txt $O/r2f_general.txt $O/r2f_timeline.txt
tail -c 600 $O/r2f_bench.json
