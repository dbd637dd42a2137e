# decisions-only cycle of every C5-sized workload, then the headline with the matrix (tools/cycle_probe.py)
for w in c5 hostname one-region redeploy mount; do
  echo "== $w"
  OE_PROBE_WORKLOAD=$w python tools/cycle_probe.py decisions 2>&1 | grep -E "cycle |oe cycle|re-deployed" | tail -3
done
echo "== c5 matrix"; python tools/cycle_probe.py matrix 2>&1 | grep -E "cycle " | tail -2
echo "== mount matrix"; OE_PROBE_WORKLOAD=mount python tools/cycle_probe.py matrix 2>&1 | grep -E "cycle " | tail -2
