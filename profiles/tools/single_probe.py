import sys, time, zlib, os
sys.path.insert(0,'/root/repo')
import numpy as np, torch
import defl_b200 as d
from defl_b200 import api, synth
import bench
n = int(sys.argv[1]) if len(sys.argv)>1 else (256<<20)
kind = sys.argv[2] if len(sys.argv)>2 else 'hist'
if kind=='hist': raw,z = bench.single_history_stream(n)
else: raw,z = synth.single_stream(n, seed=7)
dev=torch.device('cuda',0)
t_src=torch.frombuffer(bytearray(z),dtype=torch.uint8).to(dev); t_dst=torch.zeros(len(raw),dtype=torch.uint8,device=dev)
items=np.zeros(1,dtype=d.ITEM_DTYPE); items["src"],items["srclen"],items["dst"],items["dstlen"]=t_src.data_ptr(),len(z),t_dst.data_ptr(),len(raw)
res=np.zeros(1,dtype=d.RESULT_DTYPE)
flags=d.INFL_ZLIB|d.INFL_BATCH_VERIFY_ADLER|d.INFL_BATCH_DEVICE_PTRS
for i in range(3):
    torch.cuda.synchronize(); t=time.perf_counter(); api.infl_buf_batch_raw(items,flags,res); torch.cuda.synchronize(); dt=time.perf_counter()-t
    print(kind, 'ratio %.2f'%(len(raw)/len(z)), 'ms %.2f'%(dt*1e3), 'GB/s %.1f'%(len(raw)/dt/1e9), 'code',int(res['code'][0]),'info',hex(int(res['info'][0])))
