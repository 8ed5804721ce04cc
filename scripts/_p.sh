python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extension > gpurun_out/p8.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/p8.json')); print('pcg', round(d['value'],1), d['kernels']['pcg']['avg_us'], d['config']['final_cost'], d['config']['lm_iterations_per_solve'], d['config']['pcg_iterations_per_solve'])"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py tests/test_gpu_determinism.py tests/test_gpu_lm_loop.py -q -x 2>&1 | tail -3
