mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2k_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r2k_noalloc_alone.log
tail -5 gpurun_out/r2k_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/r2k_gpu_tests.log 2>&1; echo "suite rc=$?" >> gpurun_out/r2k_gpu_tests.log
tail -8 gpurun_out/r2k_gpu_tests.log
