#!/bin/bash
# quick correctness + timing check of the line solver variants on the GPU box
timeout 120 python - <<'PY'
import numpy as np, sys
sys.path.insert(0,'.')
import cfd_b200
from oracle import oracle as orc
for shape in [(40,300),(64,4096),(130,1030)]:
    o=orc.Cavity(*shape,100.0,order=1); g=cfd_b200.CavityFVM(*shape,order="redblack"); g.init_lid_driven()
    rng=np.random.default_rng(5)
    for name,scale,base in (("u",.3,0),("v",.3,0),("p",5.,0),("den",50.,1000.),("vis",2.,10.)):
        a=getattr(o,name); a[:,:]=base+scale*(rng.random(a.shape)-.5); g.upload(name,a)
    o.assemble_u(); g.assemble_u(); o.lisolv(3,2,o.u); g.lisolv(3,2,"u")
    print(shape, g.layout, "rel err", np.abs(g.download("u")-o.u).max()/np.abs(o.u).max())
PY
echo "quick rc=$?"
