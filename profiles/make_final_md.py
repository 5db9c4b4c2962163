"""Assembles profiles/r1_final.md (and copies the raw artefacts) from what one pass of
profiles/final_run.sh left in gpurun_out/.   python profiles/make_final_md.py"""
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")


def summ(*args):
    return subprocess.run([sys.executable, os.path.join(P, "summarize.py")] + list(args),
                          capture_output=True, text=True, cwd=ROOT).stdout


def last_json(path):
    return json.loads([line for line in open(path) if line.startswith("{")][-1])


for src, dst in [("final_bench.json", "r1_final_bench.json"),
                 ("final_bench_reference.json", "r1_final_bench_reference.json"),
                 ("final_configs.json", "r1_configs.json"),
                 ("final_ablation.txt", "r1_stage_ablation.txt"),
                 ("final_launches.csv", "r1_final_launches.csv")]:
    shutil.copy(os.path.join(G, src), os.path.join(P, dst))

d = last_json(os.path.join(G, "final_bench.json"))
r = last_json(os.path.join(G, "final_bench_reference.json"))
launches = summ("launches", "gpurun_out/final_launches.csv")
kernels = summ("kernel", "gpurun_out/final_step.ncu-rep")
post = summ("kernel", "gpurun_out/final_post.ncu-rep") if os.path.exists(os.path.join(G, "final_post.ncu-rep")) else ""
step = open(os.path.join(G, "final_stepbench.txt")).read()
cfgs = open(os.path.join(G, "final_configs.txt")).read()
abl = open(os.path.join(G, "final_ablation.txt")).read()
gen_path = os.path.join(G, "final_general.txt")
general = open(gen_path).read() if os.path.exists(gen_path) else ""
ro = d["roofline"]
doc = f"""# Round 1, final kernels (cfg2 step = Fbank(80) + sentence InputNormalization, 32 x 16 s, B200)

Everything below comes from ONE pass of `profiles/final_run.sh` on a fresh B200 (`gpurun`), after the last kernel
change of the round, assembled by `profiles/make_final_md.py`; raw artefacts: `profiles/r1_final_bench.json`,
`r1_final_bench_reference.json`, `r1_final_launches.csv`, `r1_configs.json`, `r1_stage_ablation.txt`
(1 / 2 / 8 GPUs: `r1_final_bench{{,_2gpu,_8gpu}}.json`).

`bench.py` ({d['steps']} steps, 8 per CUDA graph, `speechbrain_b200.pair(fb, norm)`): **{d['ms_per_step'] * 1e3:.2f} us/step =
{d['value'] / 1e6:.2f} M audio-s/s** ({d['gpu_launches'] // d['steps']} launches per step); the same modules without `pair()`:
{d['unpaired']['ms_per_step'] * 1e3:.2f} us/step (4 launches); dominant kernel alone {ro['us_per_launch']:.2f} us = {ro['achieved']:.0f} GB/s =
**{100 * ro['frac']:.1f} % of the measured HBM peak** ({ro['peak']} GB/s); end to end from pinned host memory
{d['e2e']['value'] / 1e6:.3f} M audio-s/s (PCIe-bound), {d['e2e_pcm16']['value'] / 1e6:.2f} M with int16 PCM; the reference chain on the box's
{r['cpu_baseline']['cores']} host cores (`--impl reference`): {r['value'] / 1e3:.2f} k audio-s/s. Clocks during the timed region:
{d['clocks']['sm_mhz']} / {d['clocks']['sm_max_mhz']} MHz, throttle reasons {d['clocks']['reasons']}.

## Launch list

`python bench.py --steps 8 --warmup 3 --no-cpu-baseline` exited 0, then the same command under
`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` (cold-cache, serialised: compare shares, not
absolute times). `fbank400_v3_kernel<1, 80>` is the int16-PCM instantiation (bench.py's supplementary `e2e_pcm16`
leg), `clamp_tiles_kernel` belongs to bench.py's supplementary un-paired leg, the Fill kernel to torch allocations
outside the timed step. In-graph share of the fused kernel: {ro['us_per_launch']:.1f} / {d['ms_per_step'] * 1e3:.1f} us = {100 * ro['us_per_launch'] / (d['ms_per_step'] * 1e3):.0f} %.

{launches}
In-graph cumulative timing (`profiles/stepbench.py`, CUDA events, no ncu, 8 steps per graph, programmatic dependent
launch between the kernels; batch with lengths 0.6-1.0 and zero-padded tails like bench.py's). The first four lines
are the un-paired flow, the last one the paired flow (statistics straight after the fused kernel, the apply kernel
clamps on the fly and reads x while the statistics finish):

```
{step}```

All five BASELINE.json configs (`profiles/configs_bench.py`; un-padded randn batches, lengths 0.6-1.0; cfg2 / cfg4
paired):

```
{cfgs}```

Stage ablation of the fused kernel (`profiles/exp_v3.sh`, variants built on the GPU box, un-padded batch; the
`tile_stats=0` lines include the 2 us memset of the utterance maxima that the tile records make unnecessary;
start of the round: 41.2 full / 36.6 without stage 4 / 27.7 without stages 3-4):

```
{abl}```

## General path (Stockham FFT kernel: stand-alone STFT / ISTFT, Fbank with n_fft = 512), eager calls on the cfg2 batch

`profiles/istft_bench.py`, `profiles/nfft512_bench.py` (the `n_fft=400` line of the second script is the fused
kernel called eagerly, for scale):

```
{general}```

## ncu --set full (profiles/prof_fbank.py: one paired cfg2 step, lengths = 1)

{kernels}

## ncu --set full, post_fast (cfg3 MFCC 64 x 4 s -> [64, 401, 660], MFCC-default instantiation)

{post}
"""
open(os.path.join(P, "r1_final.md"), "w").write(doc)
print(doc[:1500])
