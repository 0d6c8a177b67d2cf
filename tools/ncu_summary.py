#!/usr/bin/env python
"""Turn an .ncu-rep (``ncu --set full``) into the per-kernel table committed under profiles/: duration, DRAM bytes read and
written, L2 bytes, tensor-pipe and issue activity, shared-memory data-pipe wavefronts (LSU and tensor-core side),
registers, achieved occupancy.  Usage: tools/ncu_summary.py gpurun_out/hot_kernels.ncu-rep > profiles/.../ncu_summary.md"""
import csv
import io
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("lts__t_sectors.sum", "l2_sectors"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_lsu_pct"),
    ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_tc_pct"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("launch__registers_per_thread", "regs"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_pct"),
    ("launch__grid_size", "grid"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    name_i = col["Kernel Name"]
    print("| kernel | dur | DRAM read | DRAM write | L2 sectors (32 B) | tensor pipe % | issue % | smem pipe LSU % | smem pipe TC % | DRAM % | regs | warps active % | grid |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    seen = {}
    for r in rows[2:]:
        name = r[name_i].split("(")[0].replace("void ", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        seen[name] = seen.get(name, 0) + 1
        if seen[name] > 1:
            continue
        cells = []
        for metric, _ in WANT:
            i = col.get(metric)
            if i is None:
                cells.append("-")
                continue
            v, u = r[i], units[i]
            try:
                f = float(v)
                v = f"{f:.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {u}".strip())
        print(f"| `{name}` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
