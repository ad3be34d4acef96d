#!/bin/bash
# default bench at several step counts: usage tools/steps_sweep.sh 5 10 20
for k in "$@"; do
  python bench.py --steps $k --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > /tmp/gs.json
  python - "$k" <<'PY'
import json, sys
j = json.load(open("/tmp/gs.json"))
print("steps", sys.argv[1], "ms/step %.3f" % j["ms_per_step"], "rtfx %.0f" % j["value"], "e2e %.0f" % j["e2e"]["value"], "e2e ms %.3f" % j["e2e"]["ms_per_step"], j["clocks"]["reasons"])
PY
done
