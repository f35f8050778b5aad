import csv,collections,io,subprocess,sys
rep=sys.argv[1]
txt=subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(txt)))
for i,r in enumerate(rows):
    if 'Source' in r and 'Instructions Executed' in r: hdr=r; start=i+1; break
si=hdr.index('Source'); ei=hdr.index('Instructions Executed')
tot=collections.Counter()
for r in rows[start:]:
    if len(r)<=ei: continue
    try: n=int(r[ei].replace(',',''))
    except: continue
    op=r[si].strip().split()
    if not op: continue
    o=op[0] if not op[0].startswith('@') else op[1]
    parts=o.split('.')
    o=parts[0]+('.'+parts[1] if parts[0] in('IMAD','LDS','STS','LDG','STG','SHFL') and len(parts)>1 else '')
    tot[o]+=n
s=sum(tot.values())
for k,v in tot.most_common(int(sys.argv[2]) if len(sys.argv)>2 else 22): print(f"{k:16s} {100*v/s:5.1f}%")
