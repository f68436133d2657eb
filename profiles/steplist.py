import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
H=rows[hdr]; ki=H.index('Kernel Name'); vi=H.index('Metric Value'); ui=H.index('Metric Unit'); gi=H.index('Grid Size'); bi=H.index('Block Size')
seq=[]
for r in rows[hdr+1:]:
    if len(r)<=vi: continue
    v=float(r[vi].replace(',','')); 
    if r[ui]=='ns': v/=1000
    seq.append((r[ki][:58],r[gi],r[bi],v))
idx=[i for i,s in enumerate(seq) if 'csr_brute_kernel<0>' in s[0]]
a=idx[-2]; b=idx[-1]
tot=0
for k,g,bl,v in seq[a:b]:
    print("%-60s %-14s %-14s %7.2f"%(k,g,bl,v)); tot+=v
print("total",tot, b-a)
