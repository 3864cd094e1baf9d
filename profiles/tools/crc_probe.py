#!/usr/bin/env python3
"""CRC-32 of one large device buffer (for ncu): usage crc_probe.py [MiB]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np, torch
import defl_b200 as d
n = (int(sys.argv[1]) if len(sys.argv) > 1 else 1024) << 20
t = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
items = np.zeros(1, dtype=d.ITEM_DTYPE); items[0] = (t.data_ptr(), n, 0, 0)
crc = np.zeros(1, dtype=np.uint32)
for _ in range(3):
    d.lib().infl_crc32_batch(items.ctypes.data, 1, int(d.INFL_BATCH_DEVICE_PTRS), crc.ctypes.data)
torch.cuda.synchronize()
print(hex(int(crc[0])))
