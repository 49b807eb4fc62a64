# second-launch granularity (HDDM_DEFER_SPAN blocks of 8192 trials per CTA): S-A at 1e4 and 1250 sets, S-B
mkdir -p gpurun_out
OUT=gpurun_out/${TAG:-span}.txt
: > $OUT
for v in ${VARIANTS:-base span1 span2 span4}; do
  if [ "$v" = base ]; then unset HDDM_B200_LIB; else export HDDM_B200_LIB=/root/repo/hddm_b200/libvar_$v.so; fi
  line="$v"
  for args in "--sets 10000" "--sets 1250" "--sets 1250 --sweep B" "--sets 10000 --sweep B"; do
    r=$(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-sub $args 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('%.4g %.3f' % (j['value'], j['roofline']['kernel_ms']))")
    line="$line | $args: $r"
  done
  echo "$line" | tee -a $OUT
done
