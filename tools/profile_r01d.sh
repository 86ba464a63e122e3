# Round-1 (end, after the 1-warp-block change of knn_corr): launch list + full capture of knn_corr.  Under gpurun, after the same
# command exited 0 without ncu.  The captures of the unchanged adjacent kernels stay those of tools/profile_r01c.sh.
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
timeout 200 $B < /dev/null > gpurun_out/b_plain.log 2>&1 || exit 1
K='regex:knn|cloud|camera|zero_sp|topk|corr|tile|track_input|unpool|lift'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_r01d.csv $B < /dev/null > gpurun_out/ncu_l.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:knn_corr_kernel -s 20 -c 1 -o gpurun_out/prof_knn_corr_r01d_panoptic $B < /dev/null > /dev/null 2>&1
ls -la gpurun_out/*r01d*
