python -m pytest tests/test_model_gpu.py -x -q 2>&1 | tail -2
for e in 1 0 1 0; do
RETR_CROSS_OUTPROJ=$e python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('outproj=$e', round(d['value']), round(d['e2e']['value']), d['phases_ms_per_step'])"
done
