N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
timeout 150 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2w_bench_n${N}.log 2>gpurun_out/r2w_bench_n${N}.err; echo "rc=$?"
python tools/bench_line.py gpurun_out/r2w_bench_n${N}.log || { tail -5 gpurun_out/r2w_bench_n${N}.err; exit 1; }
timeout 150 $TR --master-port 29706 bench.py --gpus $N --steps 10 --warmup 3 --workload dexycb > gpurun_out/r2w_bench_n${N}_dexycb.log 2>gpurun_out/r2w_bench_n${N}_dexycb.err; echo "rc=$?"
python tools/bench_line.py gpurun_out/r2w_bench_n${N}_dexycb.log
