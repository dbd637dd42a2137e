# One-GPU measurement campaign: every workload's bench line, the reference arm, the ncu launch list (tag = $1).
tag=${1:-rXX}
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench_c5.log 2>&1; tail -1 gpurun_out/${tag}_bench_c5.log > gpurun_out/${tag}_bench_c5.json
for w in c4 c5-hostname c5-mount c5-one-region c5-redeploy; do
  python bench.py --steps 10 --warmup 3 --workload $w > gpurun_out/${tag}_bench_$w.log 2>&1; tail -1 gpurun_out/${tag}_bench_$w.log > gpurun_out/${tag}_bench_$w.json
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_ref.log 2>&1; tail -1 gpurun_out/${tag}_ref.log > gpurun_out/${tag}_bench_reference_arm.json
python bench.py --steps 2 --warmup 3 --no-ncu --no-oracle-check > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-ncu --no-oracle-check > gpurun_out/${tag}_ncu.log 2>&1
for f in gpurun_out/${tag}_bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read())
e=d.get("e2e",{})
print(sys.argv[1].split("bench_")[1], "value %.3g" % d["value"], "ms/step %.3f" % d.get("ms_per_step",0), "e2e %.3f ms" % e.get("cycle_latency_ms",0),
      "dec-only", e.get("decisions_only",{}).get("cycle_latency_ms"), "steady", e.get("steady_state",{}).get("cycle_latency_ms"),
      "oracle", e.get("decisions_match_oracle"), "roof", d.get("roofline",{}).get("frac"))
PY
done
