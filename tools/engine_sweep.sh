for n in "${@:-2 3}"; do python bench.py --steps ${STEPS:-12} --warmup 3 --no-cpu-baseline --engines $n 2>/dev/null | tail -1 > /tmp/gs.json; python - "$n" <<'PY'
import json, sys
j = json.load(open("/tmp/gs.json"))
print("engines", sys.argv[1], "ms/step %.3f" % j["ms_per_step"], "rtfx %.0f" % j["value"], "e2e %.0f" % j["e2e"]["value"], "e2e ms %.3f" % j["e2e"]["ms_per_step"])
PY
done
