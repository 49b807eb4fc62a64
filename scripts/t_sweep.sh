# C4 sweep over the trial count T at P = 10000 sets on one GPU (BASELINE.json configs[3]); one bench line per T.
for T in 1000 10000 100000 1000000 10000000; do
  timeout 600 python bench.py --trials $T --steps 3 --warmup 3 --no-cpu >> gpurun_out/t_sweep.jsonl 2>> gpurun_out/t_sweep.err
done
# large-trial, few sets: the HBM-facing corner (x is 80 MB, re-read by every set)
timeout 600 python bench.py --trials 10000000 --sets 148 --steps 3 --warmup 3 --no-cpu >> gpurun_out/t_sweep.jsonl 2>> gpurun_out/t_sweep.err
