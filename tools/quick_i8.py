import torch, time, sys
sys.path.insert(0,'.')
from gpmpc_b200 import engine as E, _lib
from oracle import gp_oracle as o
dev=torch.device('cuda',0)
for S in (4,5,6):
    p=o.synthetic_problem(2048,16,8192,p=3,seed=1337)
    gb=E.fit_batch(p["X"].to(dev),p["Y"].to(dev),p["ell"],p["sf2"],p["noise"],digits=S)
    Z=p["Z"].to(dev)
    m,v,j=E.predict(gb,Z)
    m2,v2,j2=E.predict(gb,Z,var_path="f64")
    print("S",S,"var rel err i8 vs dmma",((v-v2).abs().max()/v2.abs().max()).item())
    st=o.fit_state(p["X"],p["Y"][:,0],p["ell"],p["sf2"],p["noise"])
    mr,vr,jr=o.predict(p["Z"][:512],p["X"],p["ell"],p["sf2"],st["L"],st["alpha"])
    print("   vs oracle i8",o.rel_err(v[0,:512].cpu(),vr),"dmma",o.rel_err(v2[0,:512].cpu(),vr))
    for path in ("i8","f64"):
        for _ in range(3): E.predict(gb,Z,var_path=path)
        torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True);e1=torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): E.predict(gb,Z,var_path=path)
        e1.record(); torch.cuda.synchronize()
        print("   ",path,"ms/step",e0.elapsed_time(e1)/20)
    # variance-only pieces
    lib=_lib.load()
