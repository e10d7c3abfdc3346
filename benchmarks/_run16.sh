cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_tp_gpu.py -q -x 2>&1 | tail -8 > gpurun_out/t16_tp.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/bench16_tp2.json 2> gpurun_out/bench16_tp2.err
echo done
