python -m pytest tests/test_model_gpu.py -x -q -k "bf16 or scaled or host" 2>&1 | tail -2
run() {
  echo -n "$*  "
  env "$@" python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['e2e']['value']), round(d['phases_ms_per_step']['decode'],2))"
}
run A=1
run RETR_DECODE_GROUPS=3
run RETR_DECODE_GROUPS=4
run RETR_DECODE_GROUPS=1
