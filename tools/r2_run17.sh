# A/B of library variants on one GPU: parity tests with the variant, then timings (tools/ab.py)
mkdir -p gpurun_out
V=${1:-vconst}
MVT_LIB=$PWD/build/libmvt_$V.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu > gpurun_out/r2r_tests_$V.txt 2>&1; echo "variant tests rc=$?"; tail -2 gpurun_out/r2r_tests_$V.txt
AB_WORKLOADS=panoptic,droid,dexycb timeout 900 python tools/ab.py mvtracker_b200/libmvt.so build/libmvt_$V.so mvtracker_b200/libmvt.so build/libmvt_$V.so -- --no-model --no-distributions --no-adjacent --no-verify > gpurun_out/r2r_ab_$V.txt 2>&1
cat gpurun_out/r2r_ab_$V.txt
