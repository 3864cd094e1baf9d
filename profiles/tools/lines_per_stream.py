import csv, io, os, sys
from collections import defaultdict
out=open(sys.argv[1]).read(); a=int(sys.argv[2]); b=int(sys.argv[3]); nstreams=16000
lines=defaultdict(lambda:[0,0,0]); cur=None; cols=None
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0]=="File Path": cur=os.path.basename(row[1]); continue
    if row[0]=="Function Name": continue
    if row[0]=="Line No": cols={n:i for i,n in enumerate(row)}; continue
    if cur is None or cols is None or not row[0]: continue
    try: ln=int(row[0]); wi=int(row[cols["Instructions Executed"]] or 0); ti=int(row[cols["Thread Instructions Executed"]] or 0); sm=int(row[cols["# Samples"]] or 0)
    except: continue
    L=lines[(cur,ln)]; L[0]+=wi; L[1]+=ti; L[2]+=sm
src=open('/root/repo/defl_b200/csrc/warp_decode4.cuh').read().split('\n')
tot=0
for ln in range(a,b+1):
    v=lines.get(('warp_decode4.cuh',ln))
    if v and v[0]:
        tot+=v[0]
        print('%4d %8.0f lanes %4.1f smp %6d | %s'%(ln, v[0]/nstreams, v[1]/v[0], v[2], src[ln-1][:95]))
print('sum per stream', tot/nstreams)
