for e in 0 18 0 18 60; do
RETR_CROSS_PAD_KB=$e python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('pad=$e', round(d['value']), d['phases_ms_per_step']['decode'])"
done
