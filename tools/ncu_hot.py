#!/usr/bin/env python
"""Summarise an ncu report here (no GPU): key raw metrics + hottest SASS lines with their stall reasons.
Usage: python tools/ncu_hot.py gpurun_out/prof.ncu-rep [--top 25]"""
import csv
import io
import subprocess
import sys

RAW = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__block_size", "dram__bytes_read.sum",
       "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
       "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
       "smsp__sass_l1tex_data_pipe_lsu_wavefronts_mem_shared_op_ldgsts.sum", "sm__cycles_elapsed.avg",
       "l1tex__data_pipe_lsu_wavefronts.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
       "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
       "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
       "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    for m in RAW:
        if m in hdr:
            i = hdr.index(m)
            print(f"{m:75s} {vals[i]} {units[i]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = rows[1]
    ix = {k: i for i, k in enumerate(h)}
    data = rows[2:]

    def g(r, k):
        try:
            return float(r[ix[k]])
        except Exception:
            return 0.0
    tot = sum(g(r, "# Samples") for r in data)
    print(f"total samples {int(tot)}; smem wavefronts (source) {int(sum(g(r, 'L1 Wavefronts Shared') for r in data))}")
    print("--- hottest lines: idx samples execs source [top stalls]")
    order = sorted(range(len(data)), key=lambda i: -g(data[i], "# Samples"))[:top]
    for i in order:
        r = data[i]
        st = {k: g(r, k) for k in h if k.startswith("stall_") and "Not Issued" not in k and g(r, k) > 0}
        tops = ", ".join(f"{k[6:]}={int(v)}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{i:5d} {int(g(r, '# Samples')):6d} {int(g(r, 'Instructions Executed')):9d} {r[ix['Source']][:72]:72s} [{tops}]")
    if "--dump" in sys.argv:
        with open(sys.argv[sys.argv.index("--dump") + 1], "w") as f:
            for i, r in enumerate(data):
                f.write(f"{i:5d} {int(g(r, '# Samples')):6d} {int(g(r, 'Instructions Executed')):9d} {int(g(r, 'L1 Wavefronts Shared')):9d} {r[ix['Source']][:110]}\n")


if __name__ == "__main__":
    main()
