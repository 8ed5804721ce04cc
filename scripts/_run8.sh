set -x
run() { # name workload gpus steps
  timeout $5 python -m torch.distributed.run --nnodes=1 --nproc-per-node $3 --master-addr 127.0.0.1 --master-port 2951$3 bench.py --gpus $3 --workload $2 --steps $4 --warmup 3 --no-cpu-baseline > gpurun_out/$1.json 2> gpurun_out/$1.err; echo "rc=$?"; tail -2 gpurun_out/$1.err | cut -c1-300; python scripts/show_bench.py gpurun_out/$1.json
}
free -g | head -2; nproc
run r2_bench_cfg4_n8 cfg4 8 3 420
run r2_bench_cfg5_n8 cfg5 8 3 900
