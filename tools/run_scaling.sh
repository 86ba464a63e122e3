# bash tools/run_scaling.sh N   (inside gpurun --gpus N): diag + bench at N, then the smaller world sizes on the same box
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
port=29620
timeout 300 $TR --nproc-per-node $N --master-port $port tools/diag_dist.py panoptic p2p > gpurun_out/diag_dist${N}_p2p.log 2>&1
grep -h "equals\|total=\|build" gpurun_out/diag_dist${N}_p2p.log | sed 's/rank/\nrank/g' | grep "rank 0" | head -5
for n in $N 4 2; do
  [ $n -gt $N ] && continue
  [ $n -eq $N ] || [ $n -lt $N ] || continue
  port=$((port+1))
  timeout 400 $TR --nproc-per-node $n --master-port $port bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.log 2>gpurun_out/bench_n$n.err
  tail -1 gpurun_out/bench_n$n.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('N', d['n_gpus'], 'value %.4g' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v['ms'],3) for k,v in d['kernels'].items()}, 'e2e', d.get('e2e',{}).get('value'), d.get('e2e',{}).get('ms_per_step'))"
  [ $n -eq 2 ] && break
done
