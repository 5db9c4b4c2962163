"""Assembles profiles/r2_final.md (and copies the raw artefacts) from what one pass of
profiles/r2_final_run.sh left in gpurun_out/.   python profiles/make_r2_final_md.py"""
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


def read(name):
    p = os.path.join(G, name)
    return open(p).read() if os.path.exists(p) else "(missing)\n"


for src, dst in [("r2f_bench.json", "r2_final_bench.json"),
                 ("r2f_bench_reference.json", "r2_final_bench_reference.json"),
                 ("r2f_configs.json", "r2_configs.json"),
                 ("r2f_launches.csv", "r2_final_launches.csv")]:
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, dst))

d = last_json(os.path.join(G, "r2f_bench.json"))
r = last_json(os.path.join(G, "r2f_bench_reference.json"))
ro = d["roofline"]
sass = read("r2f_sass_counts.txt") if os.path.exists(os.path.join(G, "r2f_sass_counts.txt")) else open(
    os.path.join(P, "r2_sass_counts.txt")).read()
doc = f"""# Round 2, final kernels (cfg2 step = Fbank(80) + sentence InputNormalization, 32 x 16 s, B200)

Everything below comes from ONE pass of `profiles/r2_final_run.sh` on a fresh B200 (`gpurun`), after the last kernel
change of the round, assembled by `profiles/make_r2_final_md.py`; raw artefacts: `profiles/r2_final_bench.json`,
`r2_final_bench_reference.json`, `r2_final_launches.csv`, `r2_configs.json` (2 GPUs: `r2_bench_2gpu.json`).
GPU tests of the same pass: {read('r2f_tests.txt').strip().splitlines()[-1]}

`bench.py` ({d['steps']} steps, 8 per CUDA graph, the unmodified recipe flow `normalize(compute_features(wavs), wav_lens)`,
consecutive independent batches alternating between {d['run'].get('streams', 1)} streams inside a graph):
**{d['ms_per_step'] * 1e3:.2f} us/step = {d['value'] / 1e6:.2f} M audio-s/s** ({d['gpu_launches'] // d['steps']} launches per step;
with every step on ONE stream: {(d['run'].get('one_stream') or {}).get('ms_per_step', 0) * 1e3:.2f} us/step = {(d['run'].get('one_stream') or {}).get('value', 0) / 1e6:.2f} M audio-s/s,
outputs identical: {(d['run'].get('one_stream') or {}).get('outputs_identical')}); dominant
kernel alone {ro['us_per_launch']:.2f} us = {ro['achieved']:.0f} GB/s = **{100 * ro['frac']:.1f} % of the measured HBM peak** ({ro['peak']} GB/s),
whole step {100 * ro['step_frac']:.1f} %; end to end from pinned host memory {d['e2e']['value'] / 1e6:.3f} M audio-s/s (PCIe-bound),
{d['e2e_pcm16']['value'] / 1e6:.2f} M with int16 PCM in, {d['e2e_pcm16'].get('bf16_out', {}).get('value', 0) / 1e6:.2f} M with int16 PCM in + bfloat16 features out;
cfg5 shard (128 x 30 s, global normalisation, single rank) {d['cfg5']['value'] / 1e6:.2f} M audio-s/s; the reference chain on the box's
{r['cpu_baseline']['cores']} host cores (`--impl reference`): {r['value'] / 1e3:.2f} k audio-s/s. Clocks during the timed region:
{d['clocks']['sm_mhz']} / {d['clocks']['sm_max_mhz']} MHz, throttle reasons {d['clocks']['reasons']}.

## Launch list

`python bench.py --steps 8 --warmup 3 --no-cpu-baseline` exited 0, then the same command under
`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` (cold-cache, serialised: compare shares, not
absolute times). `fbank400_v3_kernel<1, 80>` is the int16-PCM instantiation (bench.py's supplementary `e2e_pcm16`
leg), `norm_stats_tiles` / `norm_reduce` / `norm_update` / `norm_apply` belong to the cfg5 leg, the Fill kernels to
torch allocations outside the timed step. In-graph share of the fused kernel: {ro['us_per_launch']:.1f} / {d['ms_per_step'] * 1e3:.1f} us = {100 * ro['us_per_launch'] / (d['ms_per_step'] * 1e3):.0f} %.

{summ('launches', 'gpurun_out/r2f_launches.csv')}
In-graph cumulative timing (`profiles/stepbench.py`, CUDA events, no ncu, 8 steps per graph, programmatic dependent
launch between the kernels; batch with lengths 0.6-1.0 and zero-padded tails like bench.py's):

```
{read('r2f_stepbench.txt')}```

Timeline of one step (`profiles/exp_timeline.sh`: experiment build with %globaltimer stamps; us after the producer's
`griddepcontrol.wait` returned; see also `profiles/r2_timeline.txt`):

```
{read('r2f_timeline.txt')}```

All five BASELINE.json configs (`profiles/configs_bench.py`; un-padded randn batches, lengths 0.6-1.0, unmodified flow):

```
{read('r2f_configs.txt')}```

## Other paths on the cfg2 batch, eager calls

`profiles/istft_bench.py`, `profiles/nfft512_bench.py` (n_fft 512: the fused 512-point kernel `fbank512.cu`, 180 us at
the start of the round through the Stockham kernel), `profiles/gaussian_bench.py` (dense filterbanks: scalar / sparse
loops against STFT + tensor-core projection; `SBFE_MEL_MMA` = 0 never / 2 whenever it fits):

```
{read('r2f_general.txt')}```

## SASS of the shipped library (`cuobjdump -sass libsbfe.so`, instruction counts per kernel family)

{sass}
## ncu --set full, the step's kernels (profiles/prof_norm.py: lengths 0.6-1.0, zero-padded tails)

{summ('kernel', 'gpurun_out/r2f_step.ncu-rep')}

## ncu --set full, fused Fbank kernel on a batch without padding (profiles/prof_fbank.py, PROF_PADDED=0)

{summ('kernel', 'gpurun_out/r2f_fbank_unpadded.ncu-rep')}

## ncu --set full, 512-point fused kernel and tensor-core projection (profiles/prof_other.py)

{summ('kernel', 'gpurun_out/r2f_other.ncu-rep')}
"""
open(os.path.join(P, "r2_final.md"), "w").write(doc)
print(doc[:1800])
