#!/bin/bash
# One GPU-box visit: parity tests, bench line, ncu launch list, ncu full captures of the top kernels.
# usage: tools/gpu_round.sh <tag> [kernel-regex ...]
set -u
TAG=${1:-r02k}; shift || true
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -3 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
tail -c 600 $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref rc=$?"
if timeout 300 python tools/one_step.py 2 > $OUT/${TAG}_plain.log 2>&1; then
  L=$(grep -o 'launches [0-9]*' $OUT/${TAG}_plain.log | awk '{print $2}'); H=$((L/2))
  I=$(grep -o 'init [0-9]*' $OUT/${TAG}_plain.log | awk '{print $2}'); I=${I:-0}
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s $((I+H)) -c $H --csv --log-file $OUT/${TAG}_launches.csv python tools/one_step.py 2 > $OUT/${TAG}_ncu_list.log 2>&1
  echo "ncu list rc=$? (skip $((I+H)) count $H)"
  for K in "$@"; do
    N=$(echo $K | tr -c 'a-zA-Z0-9_' '_')
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-2} -f -o $OUT/${TAG}_full_$N python tools/one_step.py 2 > $OUT/${TAG}_ncu_full_$N.log 2>&1
    echo "ncu full $K rc=$?"
  done
else
  echo "plain one_step failed"; tail -5 $OUT/${TAG}_plain.log
fi
