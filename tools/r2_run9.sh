mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:shard_iter -s 3 -c 1 -o gpurun_out/prof_shard_iter_r02a python tools/diag_shard.py > gpurun_out/r2i_ncu.log 2>&1
ls -la gpurun_out/prof_shard_iter_r02a.ncu-rep
