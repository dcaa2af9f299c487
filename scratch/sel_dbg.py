import sys, numpy as np
sys.path.insert(0,'/root/repo')
from ldso_b200 import capi
from tests import util
gpu = capi.load(); orc = capi.Api('/root/repo/oracle/liboracle.so','orc_')
calib='c4'
g = util.make_ctx(gpu, calib); o = util.make_ctx(orc, calib)
img = util.frames(calib)[0][0]
g.make_images(0, img); o.make_images(0, img)
pg = po = 3
for density in (1500.0, 1500.0, 6000.0):
    for rec in (0, 1):
        pass
    mg, ng, pg2 = g.select_make_maps(0, density, potential=pg)
    mo, no, po2 = o.select_make_maps(0, density, potential=po)
    d = np.argwhere(mg != mo)
    print(density, pg, po, '->', ng, no, pg2, po2, 'ndiff', len(d))
    if len(d):
        print(d[:20], mg[mg!=mo][:20], mo[mg!=mo][:20])
        ys, xs = d[:,0], d[:,1]
        print('x range', xs.min(), xs.max(), 'y range', ys.min(), ys.max())
    pg, po = pg2, po2
# no-recursion single passes at fixed potentials
for pot in (3, 5, 8, 14):
    g2 = util.make_ctx(gpu, calib); o2 = util.make_ctx(orc, calib)
    g2.make_images(0, img); o2.make_images(0, img)
    mg, ng, _ = g2.select_make_maps(0, 1500.0, potential=pot, recursions=0)
    mo, no, _ = o2.select_make_maps(0, 1500.0, potential=pot, recursions=0)
    d = np.argwhere(mg != mo)
    print('pot', pot, ng, no, len(d), d[:10].tolist())
