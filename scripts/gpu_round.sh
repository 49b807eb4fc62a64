# One GPU round on ONE B200: parity tests, smoke, the bench lines (headline with every sub-record, S-B, the reference
# arm), then the ncu passes -- each after the same command exited 0 without ncu.  TAG names the outputs under
# gpurun_out/; the executed-work JSON bench.py reads (profiles/ncu_fp64.json) is rebuilt from the captures ON THE
# BOX and comes back as gpurun_out/${TAG}_ncu_fp64.json (copy it over profiles/ncu_fp64.json and commit).
#   TAG=r2_v1 bash scripts/gpu_round.sh            (SKIP_NCU=1: tests and bench lines only)
set -x
TAG=${TAG:-r2_v1}
O=gpurun_out
mkdir -p $O
SHA=$(python -c "import bench; print(bench.csrc_sha())")
echo "csrc_sha $SHA" > $O/${TAG}_meta.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv >> $O/${TAG}_meta.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -2 $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; tail -c 600 $O/${TAG}_bench.err
timeout 600 python bench.py --sweep B --no-cpu --no-sub > $O/${TAG}_bench_sweepB.json 2>> $O/${TAG}_bench.err
timeout 600 python bench.py --sweep B --no-cpu --no-sub --trials 20000 --sets 2000 > $O/${TAG}_bench_sweepB_small.json 2>> $O/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2>> $O/${TAG}_bench.err
for f in bench bench_sweepB bench_sweepB_small bench_reference; do python - $O/${TAG}_$f.json <<'EOF'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "roofline", {k: d.get("roofline", {}).get(k) for k in ("frac", "frac_from_stale_capture", "kernel_ms")},
          "parity", d.get("parity"), "strong", (d.get("strong") or {}).get("value"), "ppc", (d.get("ppc") or {}).get("value"),
          "mcmc", {k: (v.get("it_per_s_per_chain") if isinstance(v, dict) else v) for k, v in (d.get("mcmc") or {}).items() if k != "config"})
except Exception as e:
    print(sys.argv[1], "unreadable", e)
EOF
done
[ -n "$SKIP_NCU" ] && exit 0
# ---- ncu (one tool per call).  Launch list of a small sweep, then full captures of the top kernel.
SMALL="python bench.py --steps 1 --warmup 3 --trials 20000 --sets 2000 --no-cpu --no-sub"
HEAD="python bench.py --steps 1 --warmup 3 --no-cpu --no-sub"
timeout 300 $HEAD > $O/${TAG}_plain_head.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_headline.csv $HEAD > $O/${TAG}_ncu_l.log 2>&1
# the pdf_sv evaluations of one headline step (deterministic): the denominator of "FP64 instructions per pdf_sv"
EVALS=$(python -c "import json;d=json.loads(open('$O/${TAG}_plain_head.json').read().strip().splitlines()[-1]);print(d['roofline']['pdf_sv_per_eval']*d['config']['trials']*d['config']['sets_per_gpu'])")
# (a big sweep is k_build_tables + two launches of k_like_partial per step: first pass, then the deferred trials;
#  k_like_partial launches 2 / 3 are the first / second pass of the second warm-up step)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 2 -c 1 -o $O/${TAG}_like -f $HEAD > $O/${TAG}_ncu_f.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 3 -c 1 -o $O/${TAG}_like_second -f $HEAD > $O/${TAG}_ncu_f2.log 2>&1
if [ -f $O/${TAG}_like.ncu-rep ] && [ -f $O/${TAG}_like_second.ncu-rep ]; then
  python profiles/ncu_extract.py --key A:100000x10000 --pdf-sv-evals $EVALS --sha $SHA --source "profiles/${TAG}_ncu_details_headline.txt + _second" --out $O/${TAG}_ncu_fp64.json $O/${TAG}_like.ncu-rep $O/${TAG}_like_second.ncu-rep > $O/${TAG}_ncu_extract.log 2>&1
  for r in like like_second; do
    ncu -i $O/${TAG}_$r.ncu-rep --page details > $O/${TAG}_ncu_details_$r.txt 2>&1
    python profiles/ncu_digest.py $O/${TAG}_$r.ncu-rep > $O/${TAG}_ncu_digest_$r.txt 2>&1
  done
  ncu -i $O/${TAG}_like.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${TAG}_ncu_source_like.csv.gz
fi
timeout 300 $SMALL --sweep B > $O/${TAG}_plain_B.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 2 -c 1 -o $O/${TAG}_like_B -f $SMALL --sweep B > $O/${TAG}_ncu_fB.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_like_partial -s 3 -c 1 -o $O/${TAG}_like_B_second -f $SMALL --sweep B > $O/${TAG}_ncu_fB2.log 2>&1
if [ -f $O/${TAG}_like_B.ncu-rep ] && [ -f $O/${TAG}_like_B_second.ncu-rep ]; then
  EVB=$(python -c "import json;d=json.loads(open('$O/${TAG}_plain_B.json').read().strip().splitlines()[-1]);print(d['roofline']['pdf_sv_per_eval']*d['config']['trials']*d['config']['sets_per_gpu'])")
  python profiles/ncu_extract.py --key B:20000x2000 --pdf-sv-evals $EVB --sha $SHA --source "profiles/${TAG}_ncu_details_like_B.txt + _second" --out $O/${TAG}_ncu_fp64.json $O/${TAG}_like_B.ncu-rep $O/${TAG}_like_B_second.ncu-rep >> $O/${TAG}_ncu_extract.log 2>&1
  for r in like_B like_B_second; do
    ncu -i $O/${TAG}_$r.ncu-rep --page details > $O/${TAG}_ncu_details_$r.txt 2>&1
    python profiles/ncu_digest.py $O/${TAG}_$r.ncu-rep > $O/${TAG}_ncu_digest_$r.txt 2>&1
  done
  ncu -i $O/${TAG}_like_B_second.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${TAG}_ncu_source_like_B_second.csv.gz
  rm -f $O/${TAG}_like_B.ncu-rep $O/${TAG}_like_B_second.ncu-rep
fi
cat $O/${TAG}_ncu_extract.log
du -sh $O
