for a in none sattn cattn sattn,cattn sqkv,sout,cq,cout ff1,ff2 mlp sqkv,sout,cq,cout,ff1,ff2,mlp sattn,cattn,sqkv,sout,cq,cout,ff1,ff2,mlp; do
  echo -n "ablate=$a  "
  RETR_ABLATE=$a python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['phases_ms_per_step'])"
done
