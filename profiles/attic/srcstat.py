import csv,collections,sys,subprocess
rep, kern = sys.argv[1], sys.argv[2]
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--kernel-name','regex:'+kern],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
secs=[i for i,r in enumerate(rows) if r and r[0]=='Kernel Name']
s0=secs[0]; s1=secs[1] if len(secs)>1 else len(rows)
hdr=rows[s0+1]; data=[r for r in rows[s0+2:s1] if len(r)==len(hdr)]
si=hdr.index('# Samples'); ie=hdr.index('Instructions Executed')
cols=[c for c in hdr if c.startswith('stall_') and 'Not Issued' not in c]
tot={c:0 for c in cols}
for r in data:
    for c in cols: tot[c]+=int(r[hdr.index(c)] or 0)
print(sorted(tot.items(), key=lambda kv:-kv[1])[:6])
T=sum(int(r[ie] or 0) for r in data)
print('samples',sum(int(r[si] or 0) for r in data),'inst', T, 'static', len(data))
mix=collections.Counter()
for r in data:
    op=r[1].split()
    if not op: continue
    o=op[1] if op[0].startswith('@') else op[0]
    mix[o.split('.')[0]]+=int(r[ie] or 0)
print([(k,round(100*v/T,1)) for k,v in mix.most_common(14)])
n=int(sys.argv[3]) if len(sys.argv)>3 else 15
top=sorted(range(len(data)), key=lambda i:-int(data[i][si] or 0))[:n]
for i in sorted(top): print(i, data[i][1][:64], data[i][si], data[i][ie])
