import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
kern=None; hdrs={}; secs={}
for r in rows:
    if not r: continue
    if r[0]=="Kernel Name": kern=r[1]; continue
    if r[0]=="Address": hdrs[kern]=r; continue
    secs.setdefault(kern,[]).append(r)
cols=["stall_barrier","stall_long_sb","stall_short_sb","stall_mio","stall_wait","stall_math","stall_branch_resolving","stall_selected","stall_not_selected","stall_no_inst","stall_dispatch","stall_lg"]
for k,rs in secs.items():
    hdr=hdrs[k]
    ia=hdr.index("Warp Stall Sampling (All Samples)"); ii=hdr.index("Instructions Executed")
    ci=[hdr.index(c) for c in cols]
    ie=hdr.index("L1 Wavefronts Shared Excessive"); iw=hdr.index("L1 Wavefronts Shared")
    tot=sum(int(r[ia]) for r in rs); toti=sum(int(r[ii]) for r in rs)
    print("==",k[:50],"samples",tot,"instr",toti,"smem wavefronts",sum(int(r[iw]) for r in rs),"excess",sum(int(r[ie]) for r in rs))
    seg=0; acc=[0]*len(cols); ins=0; smp=0
    print("       smp%  ins%  "+" ".join("%7s"%c[6:13] for c in cols))
    for i,r in enumerate(rs+[None]):
        if r is None or 'BAR.' in r[1]:
            print(" seg%d %5.1f %5.1f "%(seg,100*smp/tot,100*ins/toti)+" ".join("%6.1f%%"%(100*a/tot) for a in acc)); seg+=1; acc=[0]*len(cols); ins=0; smp=0
            continue
        for j,c in enumerate(ci): acc[j]+=int(r[c])
        ins+=int(r[ii]); smp+=int(r[ia])
