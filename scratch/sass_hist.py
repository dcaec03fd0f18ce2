"""Aggregate the ncu source page (SASS) of the first kernel in /tmp/src.csv by opcode: executed warp instructions and stall samples."""
import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1] if len(sys.argv)>1 else '/tmp/src.csv')))
start=[i for i,r in enumerate(rows) if r and r[0]=='Kernel Name']
seg=rows[start[0]+1:start[1]] if len(start)>1 else rows[start[0]+1:]
hdr=seg[0]; data=seg[1:]
ci=hdr.index('Instructions Executed'); si=hdr.index('Source'); ss=hdr.index('# Samples')
tot=0; byop=collections.Counter(); samp=collections.Counter()
for r in data:
    try: n=int(r[ci]); s=int(r[ss])
    except: continue
    toks=r[si].split()
    op=toks[1] if toks[0].startswith('@') else toks[0]
    op=op.split('.')[0]
    byop[op]+=n; samp[op]+=s; tot+=n
print('total warp inst',tot, 'n sass', len(data))
ts=sum(samp.values())
for op,n in byop.most_common(45):
    print(f'{op:12s} {n:12d} {100*n/tot:6.2f}%  samples {100*samp[op]/ts:6.2f}%')
