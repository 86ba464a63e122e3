# bash tools/r2_run10.sh N [quick]  (inside gpurun --gpus N): per-iteration time of the transports on N ranks + bench at N, fused vs p2p
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
L=gpurun_out/r2j_iter_n$N.txt
: > $L
timeout 300 $TR --master-port 29701 tools/diag_iter.py panoptic fused,p2p >> $L 2>gpurun_out/r2j_err1.txt
if [ -z "$2" ]; then
MVT_SHARD_EARLY_PCT=10 timeout 300 $TR --master-port 29703 tools/diag_iter.py panoptic fused >> $L 2>gpurun_out/r2j_err3.txt
timeout 300 $TR --master-port 29704 tools/diag_iter.py dexycb fused,p2p >> $L 2>gpurun_out/r2j_err4.txt
fi
grep -a "ms per iteration" $L
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['config'].get('workload','')[:12], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), d.get('e2e',{}).get('ms_per_step'), d.get('e2e',{}).get('d2h_bytes_per_step'), 'verified', d.get('verified'))"; }
for mode in fused p2p; do
  timeout 400 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 --dist-mode $mode > gpurun_out/r2j_bench_n${N}_$mode.log 2>gpurun_out/r2j_bench_n${N}_$mode.err; summ gpurun_out/r2j_bench_n${N}_$mode.log
done
if [ $N -ge 8 ]; then
  timeout 400 $TR --master-port 29706 bench.py --gpus $N --steps 10 --warmup 3 --workload dexycb > gpurun_out/r2j_bench_n${N}_dexycb.log 2>gpurun_out/r2j_bench_n${N}_dexycb.err; summ gpurun_out/r2j_bench_n${N}_dexycb.log
fi
