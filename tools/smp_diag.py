import sys, time, random
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import induction_b200
from induction_b200.util.data_loader import JSONFileDataLoader
from induction_b200.util.synthetic import make_tables
tables = make_tables(64, 700, 40, 5000, 8, seed=0)
ld = JSONFileDataLoader.from_tables(tables)
smp = ld._device(); smp.seed(0)
for (B,N,K,Q) in [(64,5,1,5)]:
    for it in range(3):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        torch.cuda.synchronize()
        e0.record(); smp.enqueue_walk(B,N,K,Q); e1.record(); smp.enqueue_gather(B,N,K,Q); e2.record()
        torch.cuda.synchronize()
        st = smp.status.cpu().numpy()
        print("walk+trace %.3f ms gather %.3f ms | cycles %d passes %d rounds %d -> %.1f cyc/pass, clock %.0f MHz if walk=all" % (
            e0.elapsed_time(e1), e1.elapsed_time(e2), st[1]*16, st[2], st[3], st[1]*16/max(st[2],1), st[1]*16/(e0.elapsed_time(e1)*1e3)))
        if st[9]:   # timeline build (INDNET_NVCC_DEFINES=-DINDNET_SMP_TIMELINE)
            print("   cycles: job setup %d | loop head %d | chunk wait+issue %d | jumps %d | rows %d | iterations %d" % (
                st[4]*16, st[5]*16, st[6]*16, st[7]*16, st[8]*16, st[9]))
