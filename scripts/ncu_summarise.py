"""Summarise an `ncu --set full` capture (or a metrics CSV) of bench.py into the tracked files the bench line and
DESIGN.md cite:
    profiles/<prefix>_ncu_full_summary.csv   one row per kernel launch kept: duration, DRAM bytes, L2->SM bytes,
                                             tensor-pipe active, issue active, registers
    profiles/r2_ncu_traffic.json             dram read/write bytes per launch of the two 16->16 kernels (roofline.traffic)
usage: python scripts/ncu_summarise.py capture.ncu-rep profiles/r2_tc16_conv_kernels <pairs per launch>"""
import csv, io, json, subprocess, sys

rep, prefix, batch = sys.argv[1], sys.argv[2], int(sys.argv[3])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = ["Kernel Name", "launch__grid_size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread"]
cols = [c for c in cols if c in idx]
out = [",".join(f'"{c} [{units[idx[c]]}]"' if units[idx[c]] else f'"{c}"' for c in cols)]
traffic = {}


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    if "conv4d" not in name and "ncnet" not in name:
        continue
    out.append(",".join('"%s"' % r[idx[c]] for c in cols))
    grid = int(float(r[idx["launch__grid_size"]].replace(",", ""))) if "launch__grid_size" in idx else 0
    rd = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
    wr = to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
    # the 16->16 launches are the largest grids of their kernels
    if "conv4d_tc_wgrad_pmw_kernel" in name or "conv4d_tc_wgrad_pm_kernel" in name:
        key = "conv4d_wgrad"
    elif "conv4d_tc_walk16_kernel" in name or "conv4d_tc_kernel<1>" in name or "conv4d_tc_kernel<true>" in name:
        key = "conv4d_fwd"
    else:
        continue
    if key not in traffic or grid > traffic[key]["grid"]:
        traffic[key] = {"batch": batch, "grid": grid, "dram_read_bytes": rd, "dram_write_bytes": wr,
                        "source": f"ncu --set full, {rep.split('/')[-1]}"}
open(prefix + "_ncu_full_summary.csv", "w").write("\n".join(out) + "\n")
json.dump(traffic, open("profiles/r2_ncu_traffic.json", "w"), indent=1)
print(json.dumps(traffic, indent=1))
