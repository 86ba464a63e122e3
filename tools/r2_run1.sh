# Round 2, first GPU pass: GPU tests (model test verbose), smoke, default bench.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 900 python -m pytest tests/test_gpu_model.py -x -q -m gpu -s > gpurun_out/r2_model.txt 2>&1; echo "model rc=$?"
tail -30 gpurun_out/r2_model.txt
timeout 1500 python -m pytest tests -q -m gpu --deselect tests/test_gpu_model.py > gpurun_out/r2_pytest.txt 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/r2_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2_smoke.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2_bench1.err
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/r2_bench1.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','verified','max_corr_err','neighbour_set_mismatches') if k in d})
    print({k:(round(v['ms'],4), round(v.get('frac',0),3)) for k,v in d['kernels'].items()})
    print('e2e', d.get('e2e'))
    print('fps', json.dumps(d.get('e2e_frames_per_s'))[:1500])
    print('cpu', d.get('cpu_baseline'))
except Exception as e:
    print('parse failed', e)
P
