O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "finish or tile_records or in_place" 2>&1 | tail -15
timeout 300 python profiles/stepbench.py 2>&1 | tail -8
SBFE_NO_INLAUNCH_FINISH=1 timeout 300 python profiles/stepbench.py 2>&1 | tail -8
