python -m pytest tests/test_train_gpu.py -x -q 2>&1 | tail -2
python bench_train.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('train', round(d['value']), d['ms_per_step'], d['loss_first_last'])"
