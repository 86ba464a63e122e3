# Round-2 evidence run (one GPU, under gpurun): launch list + ncu --set full captures of the dominant kernels (fp32 and bf16) +
# compute-sanitizer memcheck / racecheck of the fused kernels.  Each ncu command runs only after the same command exited 0 alone.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-model --no-distributions --no-adjacent --no-verify"
timeout 300 $B < /dev/null > gpurun_out/r02_plain.log 2>&1 || { tail -5 gpurun_out/r02_plain.log; exit 1; }
K='regex:knn|cloud|camera|zero_sp|topk|corr|tile|track_input|unpool|lift|shard'
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_r02.csv $B < /dev/null > gpurun_out/r02_ncu_l.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:knn_corr_kernel -s 20 -c 1 -o gpurun_out/prof_knn_corr_r02_panoptic $B < /dev/null > /dev/null 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:cloud_build_kernel -s 4 -c 1 -o gpurun_out/prof_cloud_build_r02_panoptic $B < /dev/null > /dev/null 2>&1
timeout 300 $B --dtype bf16 < /dev/null > gpurun_out/r02_plain_bf16.log 2>&1 && {
timeout 400 ncu --set full --clock-control none --import-source on -k regex:knn_corr_kernel -s 20 -c 1 -o gpurun_out/prof_knn_corr_r02_panoptic_bf16 $B --dtype bf16 < /dev/null > /dev/null 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:cloud_build_kernel -s 4 -c 1 -o gpurun_out/prof_cloud_build_r02_panoptic_bf16 $B --dtype bf16 < /dev/null > /dev/null 2>&1
}
# sanitizer: the fused kernels on a small window (tests that exercise deferral, warm start, negative depths, the frame cache)
T="python -m pytest -x -q -m gpu tests/test_gpu_parity.py -k corr_deferred_free_space_queries_identical tests/test_zz_no_alloc.py"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 $T > gpurun_out/sanitizer_memcheck_r02.txt 2>&1; echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck_r02.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest -x -q -m gpu tests/test_zz_no_alloc.py -k "twice and dtype0" > gpurun_out/sanitizer_racecheck_r02.txt 2>&1; echo "racecheck rc=$?" >> gpurun_out/sanitizer_racecheck_r02.txt
tail -3 gpurun_out/sanitizer_memcheck_r02.txt gpurun_out/sanitizer_racecheck_r02.txt
ls -la gpurun_out/*r02*
