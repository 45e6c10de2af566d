timeout 300 python bench.py --workload c5 --n-total 125000000 --steps 20 --warmup 3 --shapes none --e2e-steps 1 --run-iters 5 --cpu-seconds 1 $SHARD_EXTRA > gpurun_out/bs.json 2>gpurun_out/bs.err; echo rc=$?
python -c "
import json
l=json.loads(open('gpurun_out/bs.json').read().strip().splitlines()[-1])
print(l['ms_per_step'], l['roofline']['kernel_ms'], l['roofline']['frac'], l['first_pass_ms'], l['parity_check']['mismatches'])
print(l['step_ms'])
"
