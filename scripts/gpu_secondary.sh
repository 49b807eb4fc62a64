# ncu --set full captures of the kernels OTHER than the headline's: one launch each, from
# the workloads of bench_paths.py, each after the same command exited 0 without ncu.  Digest (headline metrics +
# SASS opcode mix, profiles/ncu_digest.py) and the details page come back under gpurun_out/${TAG}_*; the .ncu-rep
# files are deleted on the box (size).     TAG=r2_sec bash scripts/gpu_secondary.sh
set -x
TAG=${TAG:-r2_sec}
O=gpurun_out
mkdir -p $O
python bench_paths.py --reps 2 > $O/${TAG}_bench_paths.jsonl 2> $O/${TAG}_bench_paths.err; echo "bench_paths rc=$?"
cat $O/${TAG}_bench_paths.jsonl | cut -c1-220

# cap <name> <kernel regex> <launches to skip> <bench_paths workload>
cap() {
  name=$1; regex=$2; skip=$3; only=$4
  timeout 600 python bench_paths.py --reps 1 --only $only > $O/${TAG}_plain_$name.log 2>&1 || { echo "$name: plain run failed"; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -o $O/${TAG}_$name -f \
      python bench_paths.py --reps 1 --only $only > $O/${TAG}_ncu_$name.log 2>&1
  if [ -f $O/${TAG}_$name.ncu-rep ]; then
    ncu -i $O/${TAG}_$name.ncu-rep --page details > $O/${TAG}_ncu_details_$name.txt 2>&1
    python profiles/ncu_digest.py $O/${TAG}_$name.ncu-rep > $O/${TAG}_ncu_digest_$name.txt 2>&1
    rm -f $O/${TAG}_$name.ncu-rep
    head -12 $O/${TAG}_ncu_digest_$name.txt
  else
    echo "$name: no report"; tail -3 $O/${TAG}_ncu_$name.log
  fi
}
# (skip counts: the warm-up call of best_of launches the kernel once before the timed one)
cap cdf            k_cdf_array          1 cdf
cap ppc_tables     k_pdf_array          1 ppc
cap ppc_cumsum     k_cumsum_batch       1 ppc
cap ppc_sample     k_sample_batch       1 ppc
cap ppc_stats      k_ppc_stats          1 ppc
cap rl_scan        k_rl_scan            1 rlddm
cap rl_like        k_like_partial       0 rlddm
cap simulator      k_simulate_drift     1 simulator
cap simple         k_like_partial       2 simple
cap regression_v   k_like_partial       2 regression_resident
cap nodes          k_like_warp          1 nodes
cap build_warp     k_build_tables_warp  1 nodes
du -sh $O
