import csv, sys, re, collections
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=rows[2:]
ia=hdr.index("Instructions Executed"); isrc=hdr.index("Source"); ist=hdr.index("Warp Stall Sampling (All Samples)")
ncell=float(sys.argv[2]) if len(sys.argv)>2 else 9999392
tot=collections.Counter(); stall=collections.Counter()
total=0; tstall=0
lines=[]
for r in data:
    try: n=int(r[ia]); s=int(r[ist])
    except: continue
    m=re.match(r"\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[isrc])
    op=m.group(2) if m else "?"
    tot[op]+=n; stall[op]+=s; total+=n; tstall+=s
    lines.append((n,s,r[isrc].strip()))
print("total warp inst", total, "per cell (thread inst)", total*32/ncell, "stall samples", tstall)
for op,n in tot.most_common(28):
    print("%-8s %10d  %6.1f/cell  stall %5.1f%%"%(op,n,n*32/ncell,100*stall[op]/max(tstall,1)))
if len(sys.argv)>3:
    print("--- top stall lines")
    for i,(n,s,src) in enumerate(lines):
        if s>float(sys.argv[3])*tstall/100: print(i, n, "%.1f%%"%(100*s/tstall), src)
