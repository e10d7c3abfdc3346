cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemv_fused.py tests/test_llama_decode.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/t38.log 2>&1
for v in 1 2 3; do
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline >> gpurun_out/bench38.json 2>> gpurun_out/bench38.err
done
echo done
