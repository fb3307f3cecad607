#!/bin/bash
# usage: bash tools/sweep.sh "<env assignments ...>" ... : runs the short C5 bench once per argument with that environment
python -m pytest tests/test_render.py -x -q -m gpu 2>&1 | tail -3
for e in "$@"; do
  env $e python bench.py --steps 3 --warmup 2 --spp-per-step ${SPP:-16} --no-cpu-baseline --no-lens-grad 2>/dev/null | tail -1 > /tmp/line.json
  python - "$e" <<'PY'
import json, sys
d = json.load(open('/tmp/line.json'))
print(sys.argv[1], "value %.1f e2e %.1f" % (d["value"], d["e2e"]["value"]), {k: d["stats"][k] for k in ("rays", "bvh_nodes", "triangles", "lens_misses")})
PY
done
