# bash tools/r2_run16.sh N : frames mode with / without the prefetch pipeline (tests at N = 2)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
if [ $N -le 2 ]; then
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_torchrun.py -x -q -m gpu -k "frame or transports" > gpurun_out/r2q_tests.txt 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2q_tests.txt
fi
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['config'].get('workload','')[:12], d['config'].get('pipeline','')[:6], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), 'verified', d.get('verified'))" || tail -5 ${1%.log}.err; }
timeout 200 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2q_bench_n${N}_prefetch.log 2>gpurun_out/r2q_bench_n${N}_prefetch.err; summ gpurun_out/r2q_bench_n${N}_prefetch.log
if [ -z "$2" ]; then
timeout 200 $TR --master-port 29706 bench.py --gpus $N --steps 10 --warmup 3 --no-prefetch --no-e2e > gpurun_out/r2q_bench_n${N}_noprefetch.log 2>gpurun_out/r2q_bench_n${N}_noprefetch.err; summ gpurun_out/r2q_bench_n${N}_noprefetch.log
fi
