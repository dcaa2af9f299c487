import sys, numpy as np
sys.path.insert(0,'/root/repo'); 
from ldso_b200 import capi
from tests import util
api = capi.Api('/root/repo/oracle/liboracle.so','orc_')
S = util.immature_scene("small")
ctx = util.make_ctx(api, "small", n_frame_slots=8)
o = util.run_immature(ctx, S)
print("eth nan", np.isnan(o["energy_th"]).sum(), "n", len(S["u"]))
for c,t in zip(o["counts"], o["traces"]):
    g = t["status"]==0
    gt = S["gt_idepth"]
    inside = (t["idepth_min"][g] <= gt[g]) & (gt[g] <= t["idepth_max"][g])
    print(c, "good inside frac", inside.mean(), "interval", np.median(t["idepth_max"][g]-t["idepth_min"][g]), "q", np.median(t["quality"][g]))
idp, ok, st = o["act"]
cand = o["cand"]
print("cand", len(cand), "ok", ok.sum())
rel = np.abs(idp[ok==1]-S["gt_idepth"][cand][ok==1])/S["gt_idepth"][cand][ok==1]
print("rel err median", np.median(rel), "p90", np.percentile(rel,90))
print(np.unique(st, return_counts=True))
