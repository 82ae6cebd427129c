import numpy as np, sys
sys.path.insert(0,'/root/repo')
from oracle import restatement as R, signals
from pyfft_b200 import fft_pwelch
for (n,navr,dc,win) in ((2**17,63,(7.0,2.5),"Hanning"),(2**18,127,(100.0,-30.0),"Hamming"),(2**16,63,(0.,0.),"SFT3F"),(2**15,15,(3.0,1.0),"Bartlett")):
    t,x,y=signals.c2_cross(n=n,dc=dc)
    x=x+np.linspace(0,1.0,n)   # drift: provisional mean from the head is off
    kw=dict(Navr=navr,windowoverlap=0.5,windowfunction=win,detrend_style=1)
    for prec,tol in (("fp64",1e-12),("fp32",1e-5)):
        xs,ys=(x,y) if prec=="fp64" else (x.astype(np.float32),y.astype(np.float32))
        ref=R.fft_pwelch(t,xs.astype(np.float64),ys.astype(np.float64),**kw)
        for tp in (False,True):
            f,Pxy,Pxx,Pyy,Cxy,phi,info=fft_pwelch(t,xs,ys,plotit=False,precision=prec,two_pass_detrend=tp,**kw)
            e=[np.max(np.abs(a-ref[k]))/np.max(np.abs(ref[k])) for a,k in ((Pxx,"Pxx"),(Pyy,"Pyy"),(Pxy,"Pxy"))]
            print(win,n,prec,"two_pass" if tp else "fused",info.nwins,["%.1e"%v for v in e],"Cxy %.1e"%np.max(np.abs(Cxy-ref["Cxy"])), "OK" if max(e)<=tol else "FAIL")
