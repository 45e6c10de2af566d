timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -2
timeout 500 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo bench rc=$?
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench1_ref.json 2> gpurun_out/bench1_ref.err; echo ref rc=$?
