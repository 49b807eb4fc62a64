# ncu --set full capture of ONE launch of one kernel from a bench_paths.py workload:
#   bash scripts/gpu_cap.sh <name> <kernel regex> <launches to skip> <bench_paths workload>     (TAG names the outputs)
TAG=${TAG:-r2_sec}
O=gpurun_out
mkdir -p $O
name=$1; regex=$2; skip=$3; only=$4
timeout 600 python bench_paths.py --reps 1 --only $only > $O/${TAG}_plain_$name.log 2>&1 || { echo "$name: plain run failed"; exit 0; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -o $O/${TAG}_$name -f \
    python bench_paths.py --reps 1 --only $only > $O/${TAG}_ncu_$name.log 2>&1
if [ -f $O/${TAG}_$name.ncu-rep ]; then
  ncu -i $O/${TAG}_$name.ncu-rep --page details > $O/${TAG}_ncu_details_$name.txt 2>&1
  python profiles/ncu_digest.py $O/${TAG}_$name.ncu-rep > $O/${TAG}_ncu_digest_$name.txt 2>&1
  [ -n "$KEEP_REP" ] || rm -f $O/${TAG}_$name.ncu-rep
  head -40 $O/${TAG}_ncu_digest_$name.txt
else
  echo "$name: no report"; tail -3 $O/${TAG}_ncu_$name.log
fi
