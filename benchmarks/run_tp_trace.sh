#!/bin/bash
# benchmarks/mega_trace.py under torchrun: benchmarks/run_tp_trace.sh <workload> <N> <out.json> [extra args]
wl=$1; n=$2; out=$3; shift 3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
  benchmarks/mega_trace.py --workload $wl --out $out "$@" 2> ${out%.json}.err | grep -v NCCL
