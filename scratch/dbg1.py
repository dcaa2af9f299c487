import numpy as np, sys
sys.path.insert(0,'.')
from ldso_b200 import capi, build
from tests import util
np.set_printoptions(linewidth=200, precision=6)
gpu=capi.load(); orc=capi.Api(build.ORACLE_LIB,"orc_")
g=util.make_ctx(gpu,"c1"); o=util.make_ctx(orc,"c1")
gt=util.setup_tracker(g,"c1",2000); util.setup_tracker(o,"c1",2000)
I=np.concatenate([np.eye(3).reshape(-1),np.zeros(3)])
for lvl in (0,1):
    rg=g.tracker_calc_res(0,1,lvl,I,0,0,20.0); ro=o.tracker_calc_res(0,1,lvl,I,0,0,20.0)
    print(rg, ro)
    Hg,bg=g.tracker_calc_gs(0,lvl,I,0,0); Ho,bo=o.tracker_calc_gs(0,lvl,I,0,0)
    print((Hg-Ho)/np.abs(Ho).clip(1e-9)); print(Hg[6:,6:], Ho[6:,6:]); print(bg,bo)
