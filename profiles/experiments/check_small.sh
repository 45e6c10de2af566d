timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_at_size.py -m gpu -x -q -k "not tensor" 2>&1 | tail -2
bash profiles/experiments/shard_bench.sh
timeout 300 python bench.py --steps 20 --warmup 3 --shapes none --e2e-steps 1 --run-iters 5 --cpu-seconds 1 > gpurun_out/bm.json 2>gpurun_out/bm.err
python -c "
import json
l=json.loads(open('gpurun_out/bm.json').read().strip().splitlines()[-1])
print(l['ms_per_step'], l['roofline']['kernel_ms'], l['roofline']['frac'], l['first_pass_ms'], l['full_mode_ms_per_step'], l['parity_check']['mismatches'])
"
