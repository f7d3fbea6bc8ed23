set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py 2>&1 | tail -1 > gpurun_out/final_bench.json; cat gpurun_out/final_bench.json
python bench.py --impl reference --steps 1 --warmup 0 2>&1 | tail -1 > gpurun_out/final_ref.json; cat gpurun_out/final_ref.json
python bench_train.py 2>&1 | tail -1 > gpurun_out/final_train.json; cat gpurun_out/final_train.json
