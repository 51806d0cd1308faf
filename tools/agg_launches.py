import csv, collections, re, sys
with open(sys.argv[1]) as f:
    lines=[l for l in f if not l.startswith('==')]
r=csv.DictReader(lines)
agg=collections.defaultdict(lambda:[0,0.0])
tot=0
for row in r:
    name=re.sub(r'\(.*','',row['Kernel Name'])[:70]
    v=float(row['Metric Value'].replace(',',''))
    unit=row['Metric Unit']
    if unit=='ns': v/=1e3
    elif unit=='ms': v*=1e3
    agg[name][0]+=1; agg[name][1]+=v; tot+=v
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:16]:
    print(f"{k:70s} {n:6d} {t/1e3:9.3f} ms {100*t/tot:5.1f}%")
print('total ms', tot/1e3)
