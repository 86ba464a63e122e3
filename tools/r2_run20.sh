N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
timeout 200 $TR --master-port 29705 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2v_bench_n${N}.log 2>gpurun_out/r2v_bench_n${N}.err; echo "rc=$?"
python tools/bench_line.py gpurun_out/r2v_bench_n${N}.log || tail -5 gpurun_out/r2v_bench_n${N}.err
timeout 200 $TR --master-port 29706 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/r2v_ref_n${N}.log 2>gpurun_out/r2v_ref_n${N}.err; echo "ref rc=$?"; tail -c 400 gpurun_out/r2v_ref_n${N}.log
