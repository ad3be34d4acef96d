#!/bin/bash
# pipelined bench at several decoder group sizes (clips per cluster): usage tools/group_sweep.sh 16 13 11
for g in "$@"; do
  python bench.py --steps 12 --warmup 3 --no-cpu-baseline --dec-group $g 2>/dev/null | tail -1 > /tmp/gs.json
  python - "$g" <<'PY'
import json, sys
j = json.load(open("/tmp/gs.json"))
print("group", sys.argv[1], "ms/step %.3f" % j["ms_per_step"], "rtfx %.0f" % j["value"], "e2e %.0f" % j["e2e"]["value"])
PY
done
