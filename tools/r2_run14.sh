mkdir -p gpurun_out
timeout 200 python tools/diag_host.py 8 > gpurun_out/r2n_host.txt 2>&1; timeout 200 python tools/diag_host.py 2 >> gpurun_out/r2n_host.txt 2>&1; tail -3 gpurun_out/r2n_host.txt
bash tools/profile_r02.sh
