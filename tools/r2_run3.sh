set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu -s > gpurun_out/r2c_pytest.txt 2>&1; echo "pytest rc=$?"; grep -a "stock init\|filter_invalid\|lively\|inference_mode\|bf16 autocast\|passed\|failed\|Error\|error" gpurun_out/r2c_pytest.txt | head -30
for v in default inline hard12 hard16; do
  L=""; [ $v != default ] && L="MVT_LIB=$PWD/build/libmvt_$v.so"
  env $L timeout 300 python tools/diag_static.py 2>&1 | grep -a "one kernel"
done
MVT_LIB=$PWD/build/libmvt_inline.so timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:knn_corr -c 40 --csv --log-file gpurun_out/r2c_static.csv python tools/diag_static.py > /dev/null 2>&1
grep -a "knn_corr" gpurun_out/r2c_static.csv | cut -d, -f5,13- | tail -24
