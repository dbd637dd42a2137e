# decisions-only (and matrix) cycle with the commit's first pass after chunk $pc (OE_PREFIX_CHUNK) against the default
for w in c5 hostname one-region mount; do
  for pc in 1 2; do
    echo "== $w OE_PREFIX_CHUNK=$pc"
    OE_PROBE_WORKLOAD=$w OE_PREFIX_CHUNK=$pc python tools/cycle_probe.py decisions 2>&1 | grep -E "cycle " | tail -2
  done
done
for pc in 1 2; do echo "== c5 matrix pc=$pc"; OE_PREFIX_CHUNK=$pc python tools/cycle_probe.py matrix 2>&1 | grep -E "cycle " | tail -2; done
