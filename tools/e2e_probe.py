#!/usr/bin/env python3
"""Development aid (GPU box): where the end-to-end time of a 512-frame host batch goes — H2D alone, D2H alone, the
kernels of 8 x 64-frame chunks back to back, and the pipelined call."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth

B, h, w, nf = 512, 376, 1241, 2000
ex = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h, max_batch=64)
cap = ex.default_cap()
frames = torch.from_numpy(synth.frame_batch("kitti", B, pool=32)).pin_memory()
d_frames = torch.empty_like(frames, device="cuda")
d_kps = torch.empty((B, cap, 7), dtype=torch.int32, device="cuda")
d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
d_cnt = torch.empty(B, dtype=torch.int32, device="cuda")
h_kps = torch.empty((B, cap, 7), dtype=torch.int32).pin_memory()
h_desc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
h_cnt = torch.empty(B, dtype=torch.int32).pin_memory()
st = torch.cuda.Stream(); torch.cuda.set_stream(st)

def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

ms = timed(lambda: d_frames.copy_(frames, non_blocking=True))
print(f"H2D {frames.numel()/1e6:.0f} MB: {ms:.2f} ms = {frames.numel()/ms/1e6:.1f} GB/s")
def d2h():
    h_kps.copy_(d_kps, non_blocking=True); h_desc.copy_(d_desc, non_blocking=True); h_cnt.copy_(d_cnt, non_blocking=True)
ms = timed(d2h)
nb = d_kps.numel()*4 + d_desc.numel()
print(f"D2H {nb/1e6:.0f} MB: {ms:.2f} ms = {nb/ms/1e6:.1f} GB/s")
for C in (32, 64, 128, 512):
    ex2 = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h, max_batch=C)
    def chunks():
        for c in range(0, B, C):
            ex2.extract_batch_device(d_frames[c:].data_ptr(), C, w, h, d_kps[c:].data_ptr(), d_desc[c:].data_ptr(), cap,
                                     d_cnt[c:].data_ptr(), 0, st.cuda_stream)
    print(f"kernels only, chunks of {C}: {timed(chunks):.2f} ms")
    del ex2
kp = h_kps.numpy().view(X.KP_DTYPE).reshape(B, cap)
ms = timed(lambda: ex.extract_batch(frames.numpy(), kp, h_desc.numpy(), h_cnt.numpy(), cap))
print(f"pipelined host batch: {ms:.2f} ms = {B/ms*1e3:.0f} frames/s")
