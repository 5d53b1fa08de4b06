import sys, numpy as np, cv2, torch
sys.path[:0]=['oracle','vistaf-roboskin-vision-integrated-multimodal-sensor_b200']
import synthetic as S, shape_ftp as SF, ftp_oracle as FO
ref=S.make_frame(0,reference=True)
for seed in (5,7,11,2,3):
    fr=S.make_frame(seed)
    out=SF.run_batch(ref, fr[None], chunk=1, masks=True)
    P=FO.FtpParams(apply_global_shift=False,use_ecc=False); g=FO.geometry(fr.shape[0],fr.shape[1],P)
    crop=lambda f: cv2.cvtColor(f[g["y1"]:g["y2"],g["x1"]:g["x2"]],cv2.COLOR_BGR2GRAY)
    fo=FO.ftp_pair(crop(ref),crop(fr),g,P,keep=True)
    for k in ("reliable","output_reliable","contact_dilated","contact_kept"):
        d=out[k][0].cpu().numpy().astype(bool)!=fo[k]
        if d.any():
            ys,xs=np.nonzero(d); print(seed,k,int(d.sum()),list(zip(ys[:5],xs[:5])))
    hm=out["height_mm"][0].cpu().numpy(); print(seed,"height err",float(np.nanmax(np.abs(hm-fo["height_mm"]))), "keys", [k for k in fo.keys()][:30] if seed==5 else "")
