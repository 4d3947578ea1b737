#!/bin/bash
# Does a synchronous N-GPU step lose time to "the slowest GPU" or to the collectives? Run N independent 1-GPU benches
# at the same time (one per GPU, no communication) and compare their step times with the N-GPU job's per-rank times.
N=${1:-2}
for i in $(seq 0 $((N-1))); do
  CUDA_VISIBLE_DEVICES=$i python bench.py --steps 20 --warmup 5 > gpurun_out/concurrent_n1_gpu$i.json 2> gpurun_out/concurrent_n1_gpu$i.err &
done
wait
for i in $(seq 0 $((N-1))); do python -c "import json;d=json.load(open('gpurun_out/concurrent_n1_gpu$i.json'));print('independent job on GPU $i:', round(d['ms_per_step'],3), 'ms/step', d['clocks'])"; done
