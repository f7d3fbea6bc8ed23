set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/verify_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py 2>&1 | tail -1 > gpurun_out/verify_bench.json; cat gpurun_out/verify_bench.json
