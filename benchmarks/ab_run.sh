#!/bin/bash
# A/B of library builds on ONE box: benchmarks/ab_run.sh <workload> <rounds> <name> [<name> ...]   (gpurun_ab/lib_<name>.so; "head" = the in-tree build)
wl=$1; rounds=$2; shift 2
for r in $(seq $rounds); do
  for n in "$@"; do
    if [ "$n" = head ]; then unset AO_B200_LIB; else export AO_B200_LIB=$PWD/gpurun_ab/lib_$n.so; fi
    python bench.py --workload $wl --steps 100 --warmup 5 --no-cpu-baseline --no-reference-gpu 2>/dev/null | \
      python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$n', round(d['value'],1), round(d['e2e']['value'],1), round(d['roofline']['us_per_launch'],1))"
  done
done
