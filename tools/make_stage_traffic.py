"""Rebuild profiles/ncu_stage_traffic.json from the reduced `ncu --set full` captures under profiles/ (tools/ncu_raw.py
output): for the dominant kernel of each stage, DRAM bytes read + written and the duration of one launch, taken from
the NEWEST capture that holds the kernel (the `_final` files are the shipped build).  bench.py reads the file for
`roofline_kernel.traffic`."""
import json, os, re, sys
P = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles")
# stage -> (kernel name prefix as printed by ncu, frames per launch in the capture runs, capture files newest first)
FTP, TEMP = 8, 16            # bench.py --frames 32 --distinct 8: four FTP lanes of 8 frames, two temperature lanes of 16
STAGES = {
    "F.irls_plane": ("void vftp::k_irls_accum<0>", FTP, ["r02_ncu_full_sel_irls_final.txt", "r02_ncu_full_sel_irls_morph.txt"]),
    "F.irls_quadric_contact": ("void vftp::k_irls_accum<1>", FTP, ["r02_ncu_full_sel_irls_final.txt", "r02_ncu_full_sel_irls_morph.txt"]),
    "F.irls_quadric_background": ("void vftp::k_irls_accum<1>", FTP, ["r02_ncu_full_sel_irls_final.txt", "r02_ncu_full_sel_irls_morph.txt"]),
    "T.fft_cols_peaks": ("k_temp_fft_cols_peaks", TEMP, ["r02_ncu_full_temp_fft_final.txt", "r02_ncu_full_temp_fft_features.txt"]),
    "T.fft_rows": ("k_temp_fft_rows", TEMP, ["r02_ncu_full_temp_fft_final.txt", "r02_ncu_full_temp_fft_features.txt"]),
    "T.band_ifft": ("void k_temp_ifft_rows<1>", TEMP, ["r02_ncu_full_temp_fft_final.txt", "r02_ncu_full_temp_fft_features.txt"]),
    "T.features_models_fuse": ("void k_temp_features<0>", TEMP, ["r02_ncu_full_temp_fft_features.txt"]),
    "T.illum_blur20": ("void k_blur_cols<IllumStore>", TEMP, ["r02_ncu_full_blur_final.txt", "r02_ncu_full_blur.txt"]),
    "T.final_blur": ("void k_blur_cols<FinalStore>", TEMP, ["r02_ncu_full_blur_final.txt", "r02_ncu_full_blur.txt"]),
    "F.illum_blur45": ("void k_blur_cols<vftp::IllumStoreF>", FTP, ["r02_ncu_full_blur_final.txt", "r02_ncu_full_blur.txt"]),
    "F.chamfer_frontier_assemble": ("void k_chamfer2<vftp::NotMaskZero, vftp::MaskZero>", FTP, ["r02_ncu_full_ftp_unwrap_chamfer_field.txt"]),
    "F.carrier_dft_field": ("void vftp::k_ftp_field<0>", FTP, ["r02_ncu_full_ftp_unwrap_chamfer_field.txt"]),
    "F.unwrap_unionfind": ("k_uf_flatten", FTP, ["r02_ncu_full_ftp_unwrap_chamfer_field.txt"]),
}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def launches(path):
    cur, out = None, []
    for line in open(path):
        if line.startswith("== "):
            cur = {"name": line[3:].strip()}
            out.append(cur)
        elif cur is not None:
            m = re.match(r"\s+(\S+)\s+([\d.,]+)\s*(\S+)?", line)
            if m:
                cur[m.group(1)] = float(m.group(2).replace(",", "")) * UNIT.get(m.group(3) or "", 1.0)
    return out


if __name__ == "__main__":
    res = {}
    for stage, (kern, frames, files) in STAGES.items():
        for f in files:
            p = os.path.join(P, f)
            if not os.path.exists(p):
                continue
            ls = [l for l in launches(p) if l["name"].startswith(kern) and "dram__bytes_read.sum" in l]
            if not ls:
                continue
            big = max(ls, key=lambda l: l["gpu__time_duration.sum"])         # the full-size launch (captures hold small ones too)
            res[stage] = {"kernel": kern, "frames_per_launch": frames,
                          "dram_bytes_per_launch": int(big["dram__bytes_read.sum"] + big.get("dram__bytes_write.sum", 0.0)),
                          "duration_us": round(big["gpu__time_duration.sum"], 1), "source": "profiles/" + f}
            break
        else:
            print("no capture holds", kern, file=sys.stderr)
    res["_comment"] = ("dram__bytes_read.sum + dram__bytes_write.sum and gpu__time_duration.sum of the dominant kernel of a stage, from the "
                       "round-2 ncu --set full captures (bench.py --frames 32 --distinct 8, eager launches: FTP launches carry 8 frames, "
                       "temperature launches 16), newest capture per kernel: *_final.txt = the shipped build (tools/gpu_profile_r2c.sh), "
                       "the others = captures of kernels unchanged since (tools/gpu_profile_r2.sh / _r2b.sh). Made by tools/make_stage_traffic.py")
    json.dump(res, open(os.path.join(P, "ncu_stage_traffic.json"), "w"), indent=1)
    for k, v in res.items():
        if k != "_comment":
            print(f"{k:30s} {v['dram_bytes_per_launch']/1e6:9.1f} MB {v['duration_us']:9.1f} us  {v['dram_bytes_per_launch']/v['duration_us']/1e3:7.1f} GB/s  {v['source']}")
