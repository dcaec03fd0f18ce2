"""Per-phase instruction and stall-sample breakdown of an ncu source-page CSV (phases split at the given SASS indices)."""
import csv, collections, sys
path=sys.argv[1]; bounds=[int(x) for x in sys.argv[2].split(',')]
rows=list(csv.reader(open(path)))
start=[i for i,r in enumerate(rows) if r and r[0]=='Kernel Name']
seg=rows[start[0]+1:start[1]] if len(start)>1 else rows[start[0]+1:]
hdr=seg[0]; data=seg[1:]
ci=hdr.index('Instructions Executed'); ss=hdr.index('# Samples'); si=hdr.index('Source')
sc=[(j,h) for j,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
bounds=[0]+bounds+[len(data)]
for a,b in zip(bounds[:-1],bounds[1:]):
    tot=0; samp=0; st=collections.Counter(); fp64=0
    for r in data[a:b]:
        try: n=int(r[ci])
        except: continue
        tot+=n; samp+=int(r[ss])
        t=r[si].split(); op=(t[1] if t[0].startswith('@') else t[0])
        if op.startswith(('DFMA','DMUL','DADD','DSETP')): fp64+=n
        for j,h in sc:
            try: st[h]+=int(r[j])
            except: pass
    print(f'[{a},{b}) inst {tot/1e6:7.1f}M fp64 {fp64/1e6:6.1f}M samples {samp:6d} | '+', '.join(f'{k[6:]} {v}' for k,v in st.most_common(7)))
