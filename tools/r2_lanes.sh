#!/bin/bash
# adaptive passes on k_render_lanes vs k_render (OM_NO_LANES=1), then the GPU tests
TAG=${1:-lanes}; O=gpurun_out/$TAG; mkdir -p $O
( timeout 900 python -m pytest tests -m gpu -x -q -k "adaptive or redealt or scheduler or spread" ) > $O/pytest_adaptive.log 2>&1; echo "pytest adaptive rc=$?" | tee -a $O/times.txt; tail -3 $O/pytest_adaptive.log
for p in 16 32 64 256; do
  echo "== pass_spp $p lanes" | tee -a $O/times.txt
  timeout 300 python tools/adaptive_breakdown.py --pass-spp $p 2>&1 | grep -E "^rep 1|launch ms" | tail -2 | cut -c1-560 | tee -a $O/times.txt
  echo "== pass_spp $p tiles" | tee -a $O/times.txt
  OM_NO_LANES=1 timeout 300 python tools/adaptive_breakdown.py --pass-spp $p 2>&1 | grep -E "^rep 1" | tail -1 | tee -a $O/times.txt
done
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/times.txt; tail -3 $O/pytest.log
