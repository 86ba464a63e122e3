# bash tools/run_scaling.sh N   (inside gpurun --gpus N): p2p diagnostics + bench at N (strong, weak) and at N/2 (strong)
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
summ() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], d['scaling'], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e %.4g' % d.get('e2e',{}).get('value', 0), d.get('e2e',{}).get('ms_per_step'))"; }
timeout 300 $TR --nproc-per-node $N --master-port 29650 tools/diag_dist.py panoptic p2p > gpurun_out/diag_dist${N}_p2p.log 2>&1
grep -a "equals\|back-to-back\|search+idx\|build" gpurun_out/diag_dist${N}_p2p.log | sed 's/rank [1-9].*//' | grep -a "rank 0"
MVT_FULL_RECORDS=1 timeout 300 $TR --nproc-per-node $N --master-port 29651 tools/diag_dist.py panoptic p2p > gpurun_out/diag_dist${N}_p2p_full.log 2>&1
echo "-- whole 16-byte records:"; grep -a "back-to-back\|search+idx" gpurun_out/diag_dist${N}_p2p_full.log | sed 's/rank [1-9].*//' | grep -a "rank 0"
timeout 400 $TR --nproc-per-node $N --master-port 29652 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>gpurun_out/bench_n$N.err; summ gpurun_out/bench_n$N.log
timeout 400 $TR --nproc-per-node $N --master-port 29653 bench.py --gpus $N --steps 10 --warmup 3 --scaling weak > gpurun_out/bench_n${N}_weak.log 2>gpurun_out/bench_n${N}_weak.err; summ gpurun_out/bench_n${N}_weak.log
H=$((N/2))
if [ $H -ge 2 ]; then timeout 400 $TR --nproc-per-node $H --master-port 29654 bench.py --gpus $H --steps 10 --warmup 3 > gpurun_out/bench_n$H.log 2>gpurun_out/bench_n$H.err; summ gpurun_out/bench_n$H.log; fi
