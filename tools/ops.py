import json, sys
rows=json.load(open(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/ops_bf16.json'))
tot=sum(r['ms'] for r in rows)
print("total %.2f ms"%tot)
agg={}
for r in rows:
    if r['ms']>0.0:
        key=(r['kind'],r['c_in'],r['c_out'],r['k'],r['stride'],r['ileave'])
        a=agg.setdefault(key,[0,0.0,0.0,0.0])
        a[0]+=1;a[1]+=r['ms'];a[2]+=r['gflop'];a[3]+=r['out_mb']
for k,a in agg.items():
    print("kind%d ci%-4d co%-4d k%-3d s%-2d il%d  n=%d  %7.3f ms (%.3f each)  %6.1f TF/s  out %.0f MB/op"%(k+(a[0],a[1],a[1]/a[0],a[2]/a[1] if a[1] else 0,a[3]/a[0])))
