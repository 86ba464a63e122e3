mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -x -q -m gpu -k shard_iter > gpurun_out/r2e_shard.txt 2>&1; echo "shard rc=$?"; tail -25 gpurun_out/r2e_shard.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2e_pytest.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2e_pytest.txt
timeout 300 python tools/diag_static.py > gpurun_out/r2e_static.txt 2>&1; grep -a "one kernel" gpurun_out/r2e_static.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-model --no-adjacent > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/r2e_bench.json').read().strip().splitlines()[-1])
    print('ms/step %.3f'%d['ms_per_step'], {k:(round(v['ms'],4), round(v.get('frac',0),3)) for k,v in d['kernels'].items()}, d.get('verified'))
except Exception as e: print('parse failed', e)
P
tail -3 gpurun_out/r2e_bench.err
