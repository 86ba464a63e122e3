# inside gpurun --gpus 4: bench at N = 4 (frames mode) + dexycb at N = 1 and N = 4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 4"
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['config'].get('workload','')[:12], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), d.get('e2e',{}).get('ms_per_step'), d.get('e2e',{}).get('d2h_bytes_per_step'), 'verified', d.get('verified'))"; }
timeout 200 $TR --master-port 29705 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r2p_bench_n4_frames.log 2>gpurun_out/r2p_bench_n4_frames.err || { echo "panoptic failed rc=$?"; tail -5 gpurun_out/r2p_bench_n4_frames.err; exit 1; }
summ gpurun_out/r2p_bench_n4_frames.log
timeout 200 $TR --master-port 29706 bench.py --gpus 4 --steps 10 --warmup 3 --workload dexycb > gpurun_out/r2p_bench_n4_dexycb.log 2>gpurun_out/r2p_bench_n4_dexycb.err; summ gpurun_out/r2p_bench_n4_dexycb.log
timeout 200 python bench.py --steps 10 --warmup 3 --workload dexycb --no-model --no-cpu > gpurun_out/r2p_bench_n1_dexycb.log 2>gpurun_out/r2p_bench_n1_dexycb.err; summ gpurun_out/r2p_bench_n1_dexycb.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2p_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2p_smoke.log
