mkdir -p gpurun_out
: > gpurun_out/r2g_diag.txt
timeout 300 python tools/diag_shard.py >> gpurun_out/r2g_diag.txt 2>&1
for v in minb16 minb20 minb21; do MVT_LIB=$PWD/build/libmvt_$v.so timeout 200 python tools/diag_static.py 2>&1 | grep -a "one kernel" >> gpurun_out/r2g_diag.txt; done
timeout 200 python tools/diag_static.py 2>&1 | grep -a "one kernel" >> gpurun_out/r2g_diag.txt
cat gpurun_out/r2g_diag.txt | tail -30
