O=gpurun_out; mkdir -p $O
PROF_PADDED=0 ncu --set full --import-source on --clock-control none -k regex:fbank400_v3 -s 2 -c 1 -f -o $O/r2_fbank_full python profiles/prof_fbank.py 4 > /dev/null 2>&1
ls -la $O/r2_fbank_full.ncu-rep
