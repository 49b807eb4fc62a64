set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/bench_A.json 2> gpurun_out/bench_A.err
timeout 600 python bench.py --sweep B --no-cpu > gpurun_out/bench_B.json 2> gpurun_out/bench_B.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu > gpurun_out/bench_small.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_v3_launches.csv python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 3 -c 1 -o gpurun_out/r1_v3_like -f python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
