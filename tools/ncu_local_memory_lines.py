"""Source lines of an .ncu-rep that execute LOCAL-memory instructions (thread stack), by warp instructions executed:
    python tools/ncu_local_memory_lines.py gpurun_out/env_step.ncu-rep"""
import csv, subprocess, sys, collections
rep=sys.argv[1]
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],capture_output=True,text=True).stdout
rows=csv.reader(out.splitlines())
cur=None; ix=None; curline=None
L=collections.Counter(); src={}
for r in rows:
    if not r: continue
    if r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if r[0]=='Function Name': continue
    if r[0]=='Line No': ix={n:i for i,n in enumerate(r)}; continue
    if ix is None: continue
    if r[0]=='':
        sp=r[ix['Address Space']]
        if sp=='Local' and curline:
            try: L[curline]+=float(r[ix['Instructions Executed']] or 0)
            except: pass
        continue
    curline=(cur,r[0]); src[curline]=r[1].strip()[:110]
tot=sum(L.values())
print("local-memory warp instructions:", tot)
for k,v in L.most_common(40):
    print("%6.2f%%  %s:%s  %s"%(100*v/tot,k[0],k[1],src[k]))
