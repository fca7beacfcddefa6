#!/usr/bin/env python
"""ncu --set full evidence for every kernel family of the codec step that is not the convolution.

  run   (GPU box, under ncu):  ncu --set full --clock-control none -k regex:<FAMILIES> -c 400 -f -o gpurun_out/fam \
                               python tools/ncu_families.py run
        two resident cfg-2 steps (compress + decompress of the 798 223-point frame); the second is steady state.
  sum   (anywhere):            python tools/ncu_families.py sum gpurun_out/fam.ncu-rep > profiles/r2_kernel_families_ncu.md
        per kernel name the LARGEST launch: duration, DRAM bytes moved and GB/s against the measured copy peak, the
        unit percentages and which one limits.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FAMILIES = ("k_kmap_k3_blocks|k_blockmap_insert|k_sel_|k_gauss_|k_factorized_|k_octree_|k_run_flags|k_scatter_runs|"
            "k_frame_offsets|k_children|k_kmap_down2|k_hash_|k_rans_|k_lex_keys|k_ingest|k_unpack|k_pack_|k_rows_|k_gather_|k_fill|k_cap_")


def run():
    for p in (ROOT, os.path.join(ROOT, "demo-learned-point-cloud-compression_b200")):
        sys.path.insert(0, p)
    import torch
    from b200pcc import CompressionPipeline, DecompressionPipeline, synth
    from b200pcc.pipeline import load_model
    dev = torch.device("cuda:0")
    model = load_model(dev)
    frame = synth.frame_8ivfb(seed=1)
    enc = CompressionPipeline([[1.0, 1.0]], device=dev, model=model)
    dec = DecompressionPipeline(device=dev, model=model)
    staged = enc.stage_frames([frame])
    for _ in range(2):
        c, _ = enc.compress({"frames": [frame]}, staged=staged)
        dec.decompress(c[1], to_host=False)
        torch.cuda.synchronize()
    if "--capture" in sys.argv:          # the capturer's frame preparation (k_cap_*), one 640x480 measure twice
        import numpy as np
        from b200pcc.capture import FrameVoxelizer
        rng = np.random.default_rng(0)
        h, w = 480, 640
        u, v_ = np.meshgrid(np.arange(w), np.arange(h))
        z = 0.9 + 0.3 * np.sin(u / 90.0) * np.cos(v_ / 70.0)
        xyz = np.stack([(u - w / 2) * z / 500.0, (v_ - h / 2) * z / 500.0, z], -1).reshape(-1, 3).astype(np.float32)
        rgba = rng.integers(0, 256, (h * w, 4), dtype=np.uint8).view(np.float32)
        m = np.ascontiguousarray(np.concatenate([xyz, rgba], 1), dtype=np.float32)
        vox = FrameVoxelizer({"voxel_size": 0.005, "depth_clip": 1.4, "max_points": 30000}, device=dev)
        mt = torch.from_numpy(m).to(dev).reshape(-1, 4)
        for _ in range(2):
            vox.voxelize(mt)
        torch.cuda.synchronize()
    print("done")


M = {
    "dur": "gpu__time_duration.sum",
    "rd": "dram__bytes_read.sum",
    "wr": "dram__bytes_write.sum",
    "dram": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1": "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l2": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "occ": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
    "waves": "launch__waves_per_multiprocessor",
    "l2hit": "lts__t_sector_hit_rate.pct",
}
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def summarise(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ix = {k: i for i, k in enumerate(hdr)}
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]

    def val(r, key):
        i = ix.get(M[key])
        if i is None or r[i] == "":
            return float("nan")
        v = float(r[i].replace(",", ""))
        return v * SCALE.get(units[i], 1.0)

    best = {}
    count = {}
    total = {}
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]])
        name = re.sub(r"^void ", "", name)
        d = val(r, "dur")
        count[name] = count.get(name, 0) + 1
        total[name] = total.get(name, 0.0) + d
        if name not in best or d > val(best[name], "dur"):
            best[name] = r
    print("# ncu --set full, largest launch per kernel name (two resident cfg-2 steps)\n")
    print(f"Source: `{os.path.basename(rep)}` (`tools/ncu_families.py`); GB/s = (dram read + write bytes) / duration, peak = "
          f"{peak} GB/s (MEASURED_PEAKS.json).  Times under ncu are cold-cache and serialised; they rank, the bench breakdown times.\n")
    print("| kernel | launches | sum µs | largest µs | grid x block | regs | DRAM MB (r+w) | GB/s | frac of HBM peak | dram % | l1tex % | l2 % | sm % | issue % | warps active % | limiter |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for name in sorted(best, key=lambda n: -total[n]):
        r = best[name]
        d = val(r, "dur")
        b = val(r, "rd") + val(r, "wr")
        gbs = b / d / 1e3
        pct = {k: val(r, k) for k in ("dram", "l1", "l2", "sm", "issue")}
        lim = max(pct, key=lambda k: pct[k])
        lim_s = {"dram": "HBM", "l1": "L1/shared pipe", "l2": "L2", "sm": "SM pipes", "issue": "issue slots"}[lim]
        if pct[lim] < 40:
            lim_s = f"latency / launch size (max unit {pct[lim]:.0f} %: {lim_s})"
        print(f"| `{name}` | {count[name]} | {total[name]:.1f} | {d:.1f} | {int(val(r, 'grid'))} x {int(val(r, 'block'))} | {int(val(r, 'regs'))} | "
              f"{b / 1e6:.2f} | {gbs:.0f} | {gbs / peak:.3f} | {pct['dram']:.0f} | {pct['l1']:.0f} | {pct['l2']:.0f} | {pct['sm']:.0f} | "
              f"{pct['issue']:.0f} | {val(r, 'occ'):.0f} | {lim_s} |")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "run":
        run()
    elif len(sys.argv) > 2 and sys.argv[1] == "sum":
        summarise(sys.argv[2])
    elif len(sys.argv) > 1 and sys.argv[1] == "regex":
        print(FAMILIES)
    else:
        print(__doc__)
