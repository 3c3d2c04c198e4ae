"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries kept under profiles/.

  python scripts/summarize_ncu.py launches gpurun_out/X_launches.csv  profiles/X_launches_summary.txt
  python scripts/summarize_ncu.py full     gpurun_out/X.ncu-rep       profiles/X_summary.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_imma_cycles_active_realtime.avg",
    "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__inst_executed_pipe_fma.sum",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
]


def read_csv_after_header(text, first_col):
    lines = text.splitlines()
    for i, ln in enumerate(lines):
        if ln.startswith('"%s"' % first_col):
            return list(csv.DictReader(io.StringIO("\n".join(lines[i:]))))
    return []


def launches(src, dst):
    rows = read_csv_after_header(open(src).read(), "ID")
    agg = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        val *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        agg.setdefault(r["Kernel Name"][:100], []).append(val)
    tot = sum(sum(v) for v in agg.values()) or 1.0
    with open(dst, "w") as f:
        f.write("source: %s (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised)\n" % src)
        f.write("kernel | launches | total_us | mean_us | share\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            f.write("%s | %d | %.1f | %.1f | %.3f\n" % (k, len(v), sum(v) / 1e3, sum(v) / len(v) / 1e3, sum(v) / tot))
    sys.stdout.write(open(dst).read())


def full(src, dst):
    txt = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[0]
    units = rows[1]
    data = rows[2:]
    name_col = hdr.index("Kernel Name")
    with open(dst, "w") as f:
        f.write("source: %s (ncu --set full --clock-control none --import-source on)\n" % src)
        for r in data:
            f.write("== %s (ID %s)\n" % (r[name_col][:80], r[0]))
            for i, h in enumerate(hdr):
                tensorish = ("pipe_tensor" in h or "mem_tensor" in h or "pipe_tmem" in h) and h.endswith(
                    ("pct_of_peak_sustained_elapsed", "pct_of_peak_sustained_active", "_realtime.avg", ".sum")) and ".m" not in h[-40:-30]
                if h in KEYS or tensorish:          # the tensor-pipe / TMEM counters the report holds
                    f.write("  %-80s %s %s\n" % (h, r[i], units[i]))
    sys.stdout.write(open(dst).read())
    # steady-state (accumulating) launch = the one that also reads the accumulator: DRAM bytes per launch for bench.py
    def num(r, key):
        i = hdr.index(key)
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[i], 1.0)
        return float(r[i].replace(",", "")) * mult
    corr = [r for r in data if "corr_" in r[name_col]]
    if corr and len(sys.argv) > 4:
        tag = sys.argv[5] if len(sys.argv) > 5 else ""
        r = max(corr, key=lambda r: num(r, "dram__bytes_read.sum"))
        out = {"dram_bytes_per_launch": num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum"),
               "dram_bytes_read": num(r, "dram__bytes_read.sum"), "dram_bytes_write": num(r, "dram__bytes_write.sum"),
               "kernel": r[name_col][:80], "nchan": 192, "ntime_gulp": 480, "source": dst,
               "how": "ncu --set full --clock-control none, accumulating launch of `bench.py --steps 4 --warmup 3`"}
        import json
        with open(sys.argv[4], "w") as f:
            json.dump(out, f, indent=1)


def path(src, dst):
    """duration + DRAM bytes per launch (second launch of each kernel: warm instruction cache, cold data)."""
    rows = read_csv_after_header(open(src).read(), "ID")
    per = collections.OrderedDict()
    for r in rows:
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        val *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per.setdefault(r["ID"], {"name": r["Kernel Name"][:70]})[r["Metric Name"]] = val
    with open(dst, "w") as f:
        f.write("source: %s (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
                "--clock-control none; python scripts/prof_path.py)\n" % src)
        f.write("id | kernel | us | dram_read_MB | dram_write_MB | dram_GB/s\n")
        for k, m in per.items():
            us = m.get("gpu__time_duration.sum", 0) / 1e3
            rd, wr = m.get("dram__bytes_read.sum", 0) / 1e6, m.get("dram__bytes_write.sum", 0) / 1e6
            f.write("%s | %s | %.1f | %.1f | %.1f | %.0f\n" % (k, m["name"], us, rd, wr, (rd + wr) / max(us, 1e-9) * 1e3))
    sys.stdout.write(open(dst).read())


if __name__ == "__main__":
    {"launches": launches, "full": full, "path": path}[sys.argv[1]](sys.argv[2], sys.argv[3])
