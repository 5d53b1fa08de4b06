import sys, os
R=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0]=[os.path.join(R,'vistaf-roboskin-vision-integrated-multimodal-sensor_b200'),os.path.join(R,'oracle')]
import numpy as np, torch, synthetic as S, shape_ftp as SF, ftp_oracle as FO
g=S.CROP4K
ref,de=S.make_crop4k(0,reference=True),S.make_crop4k(21)
P=FO.FtpParams(apply_global_shift=False,use_ecc=False)
disc=FO.disc(g["h"],g["w"],g["cx"],g["cy"],g["r"])
geom=dict(g,circ=disc,roi=disc,apo=FO.apodization(g["h"],g["w"],g["cx"],g["cy"],g["r"],P.apod_taper))
o=FO.ftp_pair(ref,de,geom,P,keep=True)
plan=SF.FtpPlan(g["h"],g["w"],chunk=1,geom=g)
info=plan.set_reference(torch.from_numpy(ref).cuda())
print("peak",info.peak_x,info.peak_y,o["peak_f"])
out=plan.run(torch.from_numpy(de)[None].cuda(),masks=True)
rec=SF.records_to_numpy(out["records"])[0]
print("status",hex(int(rec["status"])),"residues",rec["residues"],"amp_thr",rec["amp_thr"],o["amp_thr"],"contact_thr",rec["contact_thr"],"bg",rec["bg_median"],"core",rec["core_median"])
hm=out["height_mm"][0].cpu().numpy()
d=np.abs(hm-o["height_mm"]); d[np.isnan(d)]=0
print("max",d.max(),"n>1e-3",(d>1e-3).sum(),"n>1e-4",(d>1e-4).sum(),"argmax",np.unravel_index(d.argmax(),d.shape))
for k in ("reliable","output_reliable","contact_dilated","contact_kept"):
    m=out[k][0].cpu().numpy().astype(bool); print(k,(m!=o[k]).sum(),m.sum(),o[k].sum())
for nm in ("wrapped","quality","unw","hm"):
    pl=plan.workspace_plane(nm,1).cpu().numpy().view(np.float32).reshape(g["h"],g["w"])
    ref_=dict(wrapped=o["wrapped"],quality=o["quality"],unw=o["unwrapped"],hm=o["hm"])[nm]
    rel=o["reliable"]
    dd=np.abs(pl-ref_)[rel]; print(nm,"max diff on reliable",np.nanmax(dd), "mean", np.nanmean(dd))
ys,xs=np.where(d>1e-3)
if len(ys): print("bbox of >1e-3",ys.min(),ys.max(),xs.min(),xs.max())
