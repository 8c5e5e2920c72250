"""Turn ncu outputs (brought back in gpurun_out/) into the small tracked summaries under profiles/.

  python tools/summarize_ncu.py launches gpurun_out/launches_rNN.csv profiles/rNN_launches_summary.csv "<cmd>"
  python tools/summarize_ncu.py raw gpurun_out/prof.ncu-rep profiles/rNN_<kernel>_metrics.csv
"""
import collections
import csv
import re
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "sm__pipe_tensor_cycles_active.avg.pct", "sm__warps_active.avg.pct", "launch__registers_per_thread",
        "sm__throughput.avg.pct", "lts__throughput.avg.pct", "l1tex__throughput.avg.pct", "smsp__inst_executed.sum",
        "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_tensor",
        "lts__t_bytes.sum", "smsp__cycles_active.avg", "dram__cycles_active.avg.pct")


def launches(src, dst, cmd):
    rows = list(csv.reader(open(src)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    idx = {k: i for i, k in enumerate(rows[h])}
    agg = collections.OrderedDict()
    for r in rows[h + 1:]:
        if len(r) < len(rows[h]) or r[idx["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])
        v = float(r[idx["Metric Value"]].replace(",", ""))
        unit = r[idx["Metric Unit"]]
        v = v / 1e6 if unit == "ns" else v / 1e3 if unit in ("us", "usecond") else v * 1e3 if unit in ("s", "second") else v
        n, t = agg.get(name, (0, 0.0))
        agg[name] = (n + 1, t + v)
    tot = sum(t for _, t in agg.values())
    out = [f"# ncu launch list: {cmd}", "# per-launch times are cold-cache and serialised: compare SHARES, not absolutes",
           f"# total device time in the captured window: {tot:.2f} ms", "kernel,launches,total_ms,mean_ms,share"]
    for name, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.append(f"{name},{n},{t:.3f},{t / n:.4f},{t / tot:.4f}")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


def raw(src, dst):
    txt = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    out = ["metric,unit," + ",".join(f"launch{i}" for i in range(len(rows) - 2))]
    for i, name in enumerate(hdr):
        if name == "Kernel Name" or any(k in name for k in KEEP):
            out.append(",".join([name, units[i]] + [r[i].replace(",", "") if name != "Kernel Name" else re.sub(r"\(.*", "", r[i]).replace(",", ";") for r in rows[2:]]))
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        raw(sys.argv[2], sys.argv[3])
