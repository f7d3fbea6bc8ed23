"""Join an ncu launch list (gpu__time_duration per launch) of one backbone chunk with the conv
census -> per-conv TFLOP/s and GB/s.  usage: analyze_backbone.py csv chunk_images first_row [stem_im2col launches per pass]"""
import csv, sys
def convs(blocks=(3,4,23,3), dilation=True):
    out=[("stem_im2col",None),("stem_gemm",(3*49+13,64,1,112,112)),("maxpool",None)]
    inpl, H, dil = 64, 56, 1
    for st,(planes,n) in enumerate(zip((64,128,256,512),blocks)):
        stride = 1 if st==0 else 2
        if st==3 and dilation: stride=1
        for b in range(n):
            s = stride if b==0 else 1
            Ho = H//s
            out.append((f"l{st+1}.{b}.c1",(inpl,planes,1,H,H)))
            out.append((f"l{st+1}.{b}.c2",(planes,planes,3,Ho,Ho)))
            if b==0: out.append((f"l{st+1}.{b}.ds",(inpl,planes*4,1,Ho,Ho)))
            out.append((f"l{st+1}.{b}.c3",(planes,planes*4,1,Ho,Ho)))
            inpl=planes*4; H=Ho
    return out
def main():
    path, nimg, first = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    rows=list(csv.DictReader(l for l in open(path) if not l.startswith('==')))
    # optional 4th argument: stem_im2col launches per pass (2 when both image sets ride in one pass); fold the extras
    extra=int(sys.argv[4])-1 if len(sys.argv)>4 else 0
    for _ in range(extra):
        rows[first]['Metric Value']=str(float(rows[first]['Metric Value'])+float(rows[first+1]['Metric Value'])); del rows[first+1]
    cs=convs(); tot_t=tot_f=0
    agg={}
    for i,(name,shape) in enumerate(cs):
        r=rows[first+i]; us=float(r['Metric Value'])/1000
        if shape:
            cin,cout,k,oh,ow=shape; M=nimg*oh*ow
            fl=2*M*cout*cin*k*k; by=2*(M*cin*(1 if k==1 else 1)+M*cout*(2 if name.endswith('c3') else 1))
            tf=fl/us/1e6; gb=by/us/1e3
            tot_f+=fl
        else: tf=gb=0
        tot_t+=us
        key=name.split('.')[0]+'.'+name.split('.')[-1] if '.' in name else name
        a=agg.setdefault(key,[0,0,0]); a[0]+=us; a[1]+= (fl if shape else 0); a[2]+=1
        if True or name.startswith('l3.1.') or name.startswith('l4.'):
            print(f"{name:12s} {r['Kernel Name'][21:48]:28s} {us:8.1f} us {tf:7.1f} TF/s {gb:7.0f} GB/s(min traffic)")
    print("chunk total us", tot_t, "TFLOP/s", tot_f/tot_t/1e6)
    for k,(us,fl,n) in agg.items(): print(f"{k:12s} n={n:3d} {us:9.1f} us  {fl/us/1e6 if us else 0:7.1f} TF/s  share {us/tot_t:.3f}")
if __name__ == "__main__":
    main()
