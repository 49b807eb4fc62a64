# One GPU round: parity tests, smoke, bench lines (S-A with CPU baseline, S-B, reference arm), ncu launch list and
# full captures (each after the same command exited 0 without ncu).  TAG names the outputs under gpurun_out/.
set -x
TAG=${TAG:-r1_v9}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --sweep B --no-cpu > gpurun_out/${TAG}_bench_sweepB.json 2>> gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err
timeout 300 python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu > gpurun_out/${TAG}_bench_small.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu > gpurun_out/${TAG}_ncu_l.log 2>&1
# (a big sweep is two launches of k_like_partial per step: first pass, then the deferred trials)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 4 -c 1 -o gpurun_out/${TAG}_like -f python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu_f.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 5 -c 1 -o gpurun_out/${TAG}_like_second -f python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu_f2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 4 -c 1 -o gpurun_out/${TAG}_like_B -f python bench.py --steps 1 --warmup 3 --no-cpu --sweep B --trials 20000 --sets 2000 > gpurun_out/${TAG}_ncu_fB.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 5 -c 1 -o gpurun_out/${TAG}_like_B_second -f python bench.py --steps 1 --warmup 3 --no-cpu --sweep B --trials 20000 --sets 2000 > gpurun_out/${TAG}_ncu_fB2.log 2>&1
ls -la gpurun_out | tail -20
# large-trial corner (1e7 trials, few sets): achieved HBM traffic of the likelihood kernel
timeout 900 ncu --set full --clock-control none -k regex:k_like_partial -s 4 -c 1 -o gpurun_out/${TAG}_like_T1e7 -f python bench.py --steps 1 --warmup 3 --no-cpu --trials 10000000 --sets 148 > gpurun_out/${TAG}_ncu_fT.log 2>&1
# gpurun brings back at most 64 MiB: keep the headline first-pass report, reduce the others to text on the box
for r in like_second like_B like_B_second like_T1e7; do
  f=gpurun_out/${TAG}_$r.ncu-rep
  [ -f $f ] || continue
  ncu -i $f --page details > gpurun_out/${TAG}_ncu_details_$r.txt 2>&1
  python profiles/ncu_digest.py $f > gpurun_out/${TAG}_ncu_digest_$r.txt 2>&1
  ncu -i $f --page raw --csv 2>/dev/null | gzip > gpurun_out/${TAG}_ncu_raw_$r.csv.gz
  rm -f $f
done
du -sh gpurun_out
