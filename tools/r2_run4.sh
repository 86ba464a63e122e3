mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2d_pytest.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2d_pytest.txt
: > gpurun_out/r2d_static.txt
for v in default inline hard12 hard16; do
  L=""; [ $v != default ] && L="MVT_LIB=$PWD/build/libmvt_$v.so"
  env $L timeout 300 python tools/diag_static.py >> gpurun_out/r2d_static.txt 2>&1
done
env MVT_LIB=$PWD/build/libmvt_inline.so timeout 300 python tools/diag_static.py panoptic bf16 >> gpurun_out/r2d_static.txt 2>&1
timeout 300 python tools/diag_static.py panoptic bf16 >> gpurun_out/r2d_static.txt 2>&1
grep -a "one kernel\|Error\|error" gpurun_out/r2d_static.txt
MVT_LIB=$PWD/build/libmvt_inline.so timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
