import csv,collections,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
gi=h.index('Grid Size') if 'Grid Size' in h else None
d=collections.OrderedDict()
for r in rows[1:]:
    e=d.setdefault(r[ii],{'k':r[ki],'g':r[gi] if gi is not None else ''}); e[r[mi]]=float(r[vi].replace(',',''))
ids=list(d)
last=ids[-62:]
agg=collections.OrderedDict(); tot=0
for i in last:
    k=(d[i]['k'][:52],d[i]['g']); t=d[i]['gpu__time_duration.sum']/1e3
    a=agg.setdefault(k,[0,0]); a[0]+=1; a[1]+=t; tot+=t
for k,a in agg.items(): print(f"{k[0]:52s} grid={k[1]:14s} n={a[0]:3d} us={a[1]:8.1f} avg={a[1]/a[0]:6.1f}")
print('total',round(tot,1))
