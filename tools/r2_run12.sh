# bash tools/r2_run12.sh N  (inside gpurun --gpus N): frames mode -- tests on real ranks, per-iteration time, bench vs p2p
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
if [ $N -le 2 ]; then
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_torchrun.py -x -q -m gpu -k "frame or transports" > gpurun_out/r2l_tests.txt 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2l_tests.txt
fi
L=gpurun_out/r2l_iter_n$N.txt
timeout 300 $TR --master-port 29701 tools/diag_iter.py panoptic frames,p2p > $L 2>gpurun_out/r2l_err1.txt
grep -a "ms per iteration" $L || tail -20 gpurun_out/r2l_err1.txt
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['config'].get('workload','')[:12], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), d.get('e2e',{}).get('ms_per_step'), d.get('e2e',{}).get('d2h_bytes_per_step'), 'verified', d.get('verified'))" || tail -5 ${1%.log}.err; }
for mode in frames $2; do
  timeout 400 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 --dist-mode $mode > gpurun_out/r2l_bench_n${N}_$mode.log 2>gpurun_out/r2l_bench_n${N}_$mode.err; summ gpurun_out/r2l_bench_n${N}_$mode.log
done
