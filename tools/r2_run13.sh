# bash tools/r2_run13.sh N  (inside gpurun --gpus N): bench in frames mode, panoptic then dexycb, short timeouts
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['config'].get('workload','')[:12], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), d.get('e2e',{}).get('ms_per_step'), d.get('e2e',{}).get('d2h_bytes_per_step'), 'verified', d.get('verified'))"; }
timeout 200 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2m_bench_n${N}_frames.log 2>gpurun_out/r2m_bench_n${N}_frames.err || { echo "panoptic failed rc=$?"; tail -5 gpurun_out/r2m_bench_n${N}_frames.err; exit 1; }
summ gpurun_out/r2m_bench_n${N}_frames.log
timeout 200 $TR --master-port 29706 bench.py --gpus $N --steps 10 --warmup 3 --workload dexycb > gpurun_out/r2m_bench_n${N}_dexycb.log 2>gpurun_out/r2m_bench_n${N}_dexycb.err; summ gpurun_out/r2m_bench_n${N}_dexycb.log
