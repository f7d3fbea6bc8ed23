python -m pytest tests/test_model_gpu.py tests/test_ops_gpu.py -x -q 2>&1 | tail -2
for i in 1 2; do
python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['e2e']['value']), d['phases_ms_per_step'])"
done
