mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -x -q -m gpu -k shard_iter > gpurun_out/r2h_shard.txt 2>&1; echo "shard rc=$?"; tail -5 gpurun_out/r2h_shard.txt
timeout 300 python tools/diag_shard.py > gpurun_out/r2h_diag.txt 2>&1
MVT_SHARD_NO_LOCAL=1 timeout 300 python tools/diag_shard.py >> gpurun_out/r2h_diag.txt 2>&1
cat gpurun_out/r2h_diag.txt | tail -7
