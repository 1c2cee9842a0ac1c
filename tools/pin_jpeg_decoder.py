#!/usr/bin/env python
"""Pins the oracle JPEG decoder against cv2.imdecode over shapes x images x qualities x variants (1008 cases).
Run where cv2 is importable; the CPU test suite runs a subset (tests/test_oracle_vs_cv2.py)."""
import os, sys; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, cv2
from oracle import oracle as orc
orc.build(True)
rng=np.random.default_rng(1)
def imgs(h,w):
    yield rng.integers(0,256,(h,w,3),dtype=np.uint8)
    y,x=np.mgrid[0:h,0:w]
    yield np.stack([(x*3+y)%256,(x+2*y)%256,(x*y)%256],-1).astype(np.uint8)
    a=np.full((h,w,3),255,np.uint8); a[h//4:h//2, w//4:w//2]=0; yield a
bad=0; n=0
for (h,w) in [(16,16),(17,23),(33,31),(8,8),(1,1),(3,5),(5,4),(4,7),(2,9),(9,2),(7,6),(64,48),(100,75),(15,16),(16,15),(240,319)]:
    for im in imgs(h,w):
        for q in (100,95,75,30):
            for extra in ([], [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444],[cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422],[cv2.IMWRITE_JPEG_OPTIMIZE,1],[cv2.IMWRITE_JPEG_RST_INTERVAL,3]):
                ok,enc=cv2.imencode('.jpg',im,[cv2.IMWRITE_JPEG_QUALITY,q]+extra)
                ref=cv2.imdecode(enc,cv2.IMREAD_COLOR)
                try:
                    got=orc.jpeg_decode(enc.tobytes())
                except Exception as e:
                    print('ERR',h,w,q,extra,e); bad+=1; continue
                n+=1
                if not np.array_equal(got,ref):
                    bad+=1; d=np.argwhere(got!=ref); print('MISMATCH',h,w,q,extra,len(d),d[:3], got[tuple(d[0])], ref[tuple(d[0])])
        g=cv2.cvtColor(im,cv2.COLOR_BGR2GRAY)
        ok,enc=cv2.imencode('.jpg',g,[cv2.IMWRITE_JPEG_QUALITY,90]); ref=cv2.imdecode(enc,cv2.IMREAD_COLOR); got=orc.jpeg_decode(enc.tobytes()); n+=1
        if not np.array_equal(got,ref): bad+=1; print('GRAY MISMATCH',h,w)
print('cases',n,'bad',bad)
