TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 4 --master-port 29611 tools/diag_dist.py panoptic p2p > gpurun_out/diag_dist4_p2p.log 2>&1; tail -12 gpurun_out/diag_dist4_p2p.log
timeout 300 $TR --nproc-per-node 4 --master-port 29612 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench_n4.log 2>&1; tail -2 gpurun_out/bench_n4.log | cut -c1-1500
timeout 300 $TR --nproc-per-node 2 --master-port 29613 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2>&1; tail -1 gpurun_out/bench_n2.log | cut -c1-600
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.log 2>&1; tail -1 gpurun_out/bench_n1.log | cut -c1-600
