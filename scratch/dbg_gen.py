import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
from libultrahdr_dev_b200 import abi, synth
from libultrahdr_dev_b200.api import JpegR, take_map
from oracle.pyoracle import CpuImpl
uh, orc = JpegR(), CpuImpl.oracle()
w,h=640,360
for kind in ('natural','uniform'):
  for tf in (1,2,0):
    for sg,hg in ((0,2),(2,2),(0,0)):
        if kind=='natural': s,hd = synth.natural_pair(w,h,11+tf,tf,sg,hg)
        else: s,hd = synth.uniform_pair(w,h,17+tf)
        sdr=abi.host_yuv420(s,w,h,sg); hdr=abi.host_p010(hd,w,h,hg)
        _,want,_,_=orc.generate_gainmap(sdr,hdr,tf)
        md,dest=abi.Metadata(),abi.Image(); uh.generateGainMap(sdr,hdr,tf,md,dest); got=take_map(dest)
        d=got.astype(int)-want.astype(int)
        mid=(want>0)&(want<255)
        print(kind,tf,sg,hg,'exact',(d==0).mean(),'mid frac',mid.mean(),'exact in mid',(d[mid]==0).mean() if mid.any() else None,'+1',(d==1).sum(),'-1',(d==-1).sum())
