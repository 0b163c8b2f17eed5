import csv, sys
rows=list(csv.reader(open(sys.argv[1] if len(sys.argv)>1 else '/root/repo/gpurun_out/tc_launches.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[hi]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
tot=0; out=[]
for r in rows[hi+1:]:
    v=float(r[vi].replace(',','')); v = v/1e3 if r[ui]=='ns' else v
    tot+=v; out.append(f"{r[ki][:28]}:{v:.0f}")
print(' | '.join(out)); print('total us', tot)
