mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -x -q -m gpu -k shard_iter > gpurun_out/r2h_shard.txt 2>&1; echo "shard rc=$?"; tail -5 gpurun_out/r2h_shard.txt
MVT_SHARD_EARLY_PCT=25 timeout 600 python -m pytest tests/test_gpu_shard.py -x -q -m gpu -k shard_iter 2>&1 | tail -3
: > gpurun_out/r2h_diag.txt
timeout 300 python tools/diag_shard.py >> gpurun_out/r2h_diag.txt 2>&1
cat gpurun_out/r2h_diag.txt | tail -4
timeout 600 ncu --set full --clock-control none --import-source on -k regex:shard_iter -s 3 -c 1 -o gpurun_out/prof_shard_iter_r02b python tools/diag_shard.py > gpurun_out/r2i_ncu.log 2>&1
ls -la gpurun_out/prof_shard_iter_r02b.ncu-rep
