SBFE_NVCC_EXTRA="-DSBFE_EXP_TIMELINE" python -m speechbrain_b200._build --force > /dev/null 2>&1
python profiles/exp_timeline.py
python -m speechbrain_b200._build --force > /dev/null 2>&1
