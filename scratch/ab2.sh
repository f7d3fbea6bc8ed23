for a in none cattn; do
  echo -n "ablate=$a  "
  RETR_ABLATE=$a python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['phases_ms_per_step'])"
done
