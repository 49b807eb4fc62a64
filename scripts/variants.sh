# Side-by-side measurement of kernel build variants (hddm_b200/libvar_<name>.so, built with
# HDDM_B200_LIB / HDDM_B200_NVCC_EXTRA): parity subset, then the S-A and S-B bench lines.
mkdir -p gpurun_out
OUT=gpurun_out/${TAG:-variants}.txt
: > $OUT
for v in ${VARIANTS:-base}; do
  export HDDM_B200_LIB=/root/repo/hddm_b200/libvar_$v.so
  rc=skipped
  if [ -n "$PARITY" ]; then timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/var_${v}_pytest.log 2>&1; rc=$?; fi
  a=$(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-sub 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['value'], j['ms_per_step'])")
  b=$(timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-sub --sweep B 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['value'])")
  echo "$v parity_rc=$rc S-A $a S-B $b" | tee -a $OUT
done
