set -x
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/b_plain.log 2>&1 || exit 1
K='regex:knn|cloud|camera|zero_sp|topk|corr|tile'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cloud_build -s 6 -c 1 -o gpurun_out/prof_cloud_build_r01b_panoptic python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:knn_corr_kernel -s 20 -c 1 -o gpurun_out/prof_knn_corr_r01b_droid python bench.py --workload droid --steps 2 --warmup 3 --no-cpu --no-e2e > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:knn_keys -s 20 -c 1 -o gpurun_out/prof_knn_keys_r01b_half python tools/diag_keys.py > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:corr_from_idx -s 20 -c 1 -o gpurun_out/prof_corr_from_idx_r01b_quarter python tools/diag_keys.py > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
