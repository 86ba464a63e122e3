# 2 GPUs: real-rank transports vs fused kernel, bench at N = 1 and 2 (fused / p2p)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_torchrun.py tests/test_gpu_input.py -x -q -m gpu > gpurun_out/r2f_pytest.txt 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2f_pytest.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
B="--steps 10 --warmup 3 --no-cpu --no-model --no-adjacent --no-distributions"
summ() { python - "$1" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'N', d['n_gpus'], 'value %.4g'%d['value'], 'ms/step %.3f'%d['ms_per_step'], {k:round(v['ms'],4) for k,v in d['kernels'].items()}, 'verified', d.get('verified'), d.get('neighbour_set_mismatches'), 'e2e ms', d.get('e2e',{}).get('ms_per_step'), 'd2h', d.get('e2e',{}).get('d2h_bytes_per_step'))
except Exception as e:
    print(sys.argv[1], 'parse failed', e)
P
}
timeout 600 python bench.py --gpus 1 $B > gpurun_out/r2f_n1.json 2> gpurun_out/r2f_n1.err; summ gpurun_out/r2f_n1.json; tail -3 gpurun_out/r2f_n1.err
timeout 600 $TR --nproc-per-node 2 --master-port 29660 bench.py --gpus 2 $B > gpurun_out/r2f_n2.json 2> gpurun_out/r2f_n2.err; summ gpurun_out/r2f_n2.json; tail -5 gpurun_out/r2f_n2.err
timeout 600 $TR --nproc-per-node 2 --master-port 29661 bench.py --gpus 2 $B --dist-mode p2p > gpurun_out/r2f_n2_p2p.json 2> gpurun_out/r2f_n2_p2p.err; summ gpurun_out/r2f_n2_p2p.json; tail -3 gpurun_out/r2f_n2_p2p.err
timeout 600 $TR --nproc-per-node 2 --master-port 29662 bench.py --gpus 2 $B --step window > gpurun_out/r2f_n2_win.json 2> gpurun_out/r2f_n2_win.err; summ gpurun_out/r2f_n2_win.json; tail -3 gpurun_out/r2f_n2_win.err
