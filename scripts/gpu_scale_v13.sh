N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29542 bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/bench_n${N}_v13.json 2> gpurun_out/bench_n${N}_v13.err; cut -c1-330 gpurun_out/bench_n${N}_v13.json; tail -2 gpurun_out/bench_n${N}_v13.err
timeout 300 $TR --master-port 29541 bench.py --gpus $N --workload cornell --progressive 600 > gpurun_out/bench_c5_n${N}_v13.json 2> gpurun_out/bench_c5_n${N}_v13.err; tail -c 1300 gpurun_out/bench_c5_n${N}_v13.json | cut -c1-420; tail -2 gpurun_out/bench_c5_n${N}_v13.err
