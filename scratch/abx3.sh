python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['e2e']['value']), d['phases_ms_per_step'], d['roofline']['frac'])"
python bench_train.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('train', round(d['value']), d['ms_per_step'], d['loss_first_last'])"
