#!/bin/bash
# bench.py under torchrun for one (workload, N): benchmarks/run_tp.sh <workload> <N> <steps> <out.json>
wl=$1; n=$2; steps=$3; out=$4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
  bench.py --gpus $n --workload $wl --steps $steps --warmup 5 --no-cpu-baseline --no-reference-gpu > $out 2> ${out%.json}.err
python - <<PY
import json
try:
    d = json.load(open("$out"))
    r = d.get("roofline", {})
    print("$wl N=$n", {k: round(d[k], 2) for k in ("value", "ms_per_step")}, "e2e", round(d["e2e"]["value"], 1), "kernel_us", round(r.get("us_per_launch", 0), 1), "frac", round(r.get("frac", 0), 3))
except Exception as e:
    print("$wl N=$n FAILED", e); print(open("${out%.json}.err").read()[-1500:])
PY
