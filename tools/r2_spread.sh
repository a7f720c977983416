#!/bin/bash
# adaptive passes: k_render_spread thresholds, then the GPU tests
TAG=${1:-spread}; O=gpurun_out/$TAG; mkdir -p $O
( timeout 900 python -m pytest tests -m gpu -x -q -k "adaptive or redealt or scheduler" ) > $O/pytest_adaptive.log 2>&1; echo "pytest adaptive rc=$?" | tee -a $O/times.txt; tail -3 $O/pytest_adaptive.log
for mx in 0 50000 200000 400000 1000000 3000000; do
  echo "== OM_SPREAD_MAX_LIVE=$mx" | tee -a $O/times.txt
  if [ $mx = 0 ]; then OM_NO_SPREAD=1 timeout 300 python tools/adaptive_breakdown.py 2>&1 | grep -E "^rep 1|launch ms" | tail -2 | cut -c1-700 | tee -a $O/times.txt
  else OM_SPREAD_MAX_LIVE=$mx timeout 300 python tools/adaptive_breakdown.py 2>&1 | grep -E "^rep 1|launch ms" | tail -2 | cut -c1-700 | tee -a $O/times.txt; fi
done
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/times.txt; tail -3 $O/pytest.log
