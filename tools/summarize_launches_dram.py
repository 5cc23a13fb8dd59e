import csv, collections, sys
f=sys.argv[1]
lines=open(f).read().splitlines()
i=[k for k,l in enumerate(lines) if l.startswith('"ID"')][0]
rows=list(csv.DictReader(lines[i:]))
per=collections.defaultdict(dict)
for r in rows:
    v=float(r['Metric Value'].replace(',',''))
    u=r['Metric Unit']
    if r['Metric Name']=='gpu__time_duration.sum': v = v/1000 if u in('ns','nsecond') else v
    else: v = v/1e6 if u=='byte' else (v if u=='Mbyte' else v*1e3 if u=='Gbyte' else v/1e3 if u=='Kbyte' else v)
    per[int(r['ID'])][r['Metric Name']]=v; per[int(r['ID'])]['name']=r['Kernel Name'][:70]+' '+r.get('Grid Size','')
agg=collections.defaultdict(lambda:[0,0,0,0])
for k,d in per.items():
    a=agg[d['name']]; a[0]+=1; a[1]+=d['gpu__time_duration.sum']; a[2]+=d.get('dram__bytes_read.sum',0); a[3]+=d.get('dram__bytes_write.sum',0)
tot=sum(a[1] for a in agg.values())
print('total %.1f us over %d launches; dram read %.1f MB, write %.1f MB'%(tot,len(per),sum(a[2] for a in agg.values()),sum(a[3] for a in agg.values())))
for name,a in sorted(agg.items(), key=lambda x:-x[1][1])[:14]:
    print('%9.1f us %4dx avg %7.2f us %5.1f%%  read %7.2f MB write %6.2f MB per launch  %s'%(a[1],a[0],a[1]/a[0],100*a[1]/tot,a[2]/a[0],a[3]/a[0],name))
