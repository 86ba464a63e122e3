# Round-1 (end) ncu evidence: run under gpurun AFTER the same commands exited 0 without ncu.  Outputs under gpurun_out/.
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
timeout 200 $B < /dev/null > gpurun_out/b_plain.log 2>&1 || exit 1
K='regex:knn|cloud|camera|zero_sp|topk|corr|tile|track_input|unpool|lift'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_r01c.csv $B < /dev/null > gpurun_out/ncu_l.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:knn_corr_kernel -s 20 -c 1 -o gpurun_out/prof_knn_corr_r01c_panoptic $B < /dev/null > /dev/null 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:track_input -s 2 -c 1 -o gpurun_out/prof_track_input_r01c_panoptic $B < /dev/null > /dev/null 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:corr_backward -s 2 -c 1 -o gpurun_out/prof_corr_backward_r01c_panoptic $B < /dev/null > /dev/null 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:feature_unpool -s 2 -c 1 -o gpurun_out/prof_feature_unpool_r01c_panoptic $B < /dev/null > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
