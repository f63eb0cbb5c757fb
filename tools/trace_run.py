import os, sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam_b200 import ORBextractor
from orb_slam_b200._lib import check
from orb_slam_b200.synth import synth_frame
W,H,B=1241,376,1024
fr=np.stack([synth_frame(W,H,i) for i in range(8)])
host=torch.empty((B,H,W),dtype=torch.uint8,pin_memory=True); host.numpy()[:]=np.concatenate([fr]*(B//8))
ex=ORBextractor(2000,1.2,8,20,7); L=ex._L; h=ex.handle; cap=ex.max_keypoints(W,H)
k=torch.empty((B,cap,28),dtype=torch.uint8,pin_memory=True); d=torch.empty((B,cap,32),dtype=torch.uint8,pin_memory=True); n=np.zeros(B,np.int32)
ptrs=(C.c_void_p*B)(*[host.data_ptr()+i*W*H for i in range(B)])
for r in range(3):
    print("--- call", r, file=sys.stderr)
    check(L.orbx_extract_batch(h,ptrs,B,W,H,W,k.data_ptr(),d.data_ptr(),cap,n.ctypes.data),"x")
