O=gpurun_out; mkdir -p $O
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 50 --warmup 5 > $O/r2_bench_${N}gpu.json 2> $O/r2_bench_${N}gpu.err; python - <<PY
import json
d=json.loads([l for l in open("$O/r2_bench_${N}gpu.json") if l.startswith("{")][-1])
print("N=%d value %.2f M  ms/step %.4f  e2e %.3f M  pcm16 %.3f M  pcm16+bf16 %.3f M" % (d["n_gpus"], d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["e2e_pcm16"]["value"]/1e6, d["e2e_pcm16"]["bf16_out"]["value"]/1e6))
c=d["cfg5"]; print("cfg5", c["parallelism"], "no_sync %.2f M  with_sync %.2f M  exposed allreduce %.2f us" % (c["no_sync"]["value"]/1e6, c.get("with_sync",{}).get("value",0)/1e6, c.get("exposed_allreduce_us_per_step",0)))
PY
tail -2 $O/r2_bench_${N}gpu.err
