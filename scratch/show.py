import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e); continue
    print(f, {k:round(v,2) for k,v in d.items() if isinstance(v,float)})
    for k,v in d.items():
        if isinstance(v,dict): print("   ",k,{kk:(round(vv,1) if isinstance(vv,float) else vv) for kk,vv in v.items() if not kk.startswith("phases")},{a:round(b,1) for a,b in v["phases0"].items()},{a:round(b,1) for a,b in v["phases1"].items()})
