O=gpurun_out; mkdir -p $O
python -m pytest tests/test_dist_gpu.py -m gpu -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 50 --warmup 5 > $O/r2_bench_2gpu.json 2> $O/r2_bench_2gpu.err; tail -c 1500 $O/r2_bench_2gpu.json; tail -3 $O/r2_bench_2gpu.err
