# usage: _run_n.sh GPUS WORKLOAD [more workloads]   -- bench lines of sharded workloads at one GPU count
N=$1; shift
for w in "$@"; do
  if [ "$N" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_${w}_n1.json 2> gpurun_out/r2_bench_${w}_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_${w}_n$N.json 2> gpurun_out/r2_bench_${w}_n$N.err
  fi
  echo "rc=$?"; python scripts/show_bench.py gpurun_out/r2_bench_${w}_n$N.json
done
