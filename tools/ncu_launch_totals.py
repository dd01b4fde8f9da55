import csv,collections,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
H=rows[hdr]; data=rows[hdr+1:]
ix={n:i for i,n in enumerate(H)}
agg=collections.OrderedDict()
for r in data:
    if len(r)<len(H): continue
    i=int(r[ix['ID']]); k=r[ix['Kernel Name']].split('<')[0].split('(')[0]
    agg.setdefault((i,k),{})[r[ix['Metric Name']]]=float(r[ix['Metric Value']].replace(',',''))
tot=collections.defaultdict(float); series=collections.defaultdict(list); issue=collections.defaultdict(list)
for (i,k),m in agg.items():
    tot[k]+=m['gpu__time_duration.sum']/1e6; series[k].append(round(m['gpu__time_duration.sum']/1e6,1))
    if 'smsp__issue_active.avg.pct_of_peak_sustained_active' in m: issue[k].append(round(m['smsp__issue_active.avg.pct_of_peak_sustained_active']))
for k in tot: print(k, 'total ms', round(tot[k],1)); print('  t', series[k]); print('  issue%', issue[k])
