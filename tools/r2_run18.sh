# packed bf16 cloud build: tests with the default library, then A/B against -DMVT_CLOUD_PACKED=0
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_input.py tests/test_gpu_dropin.py -x -q -m gpu -k "cloud or window_sequence or bf16 or dropin" > gpurun_out/r2s_tests.txt 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2s_tests.txt
AB_WORKLOADS=panoptic,droid timeout 600 python tools/ab.py build/libmvt_unpacked.so mvtracker_b200/libmvt.so build/libmvt_unpacked.so mvtracker_b200/libmvt.so -- --dtype bf16 --no-model --no-distributions --no-adjacent > gpurun_out/r2s_ab.txt 2>&1
cat gpurun_out/r2s_ab.txt
