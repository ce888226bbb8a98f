"""One line per profiled launch from `ncu -i report.ncu-rep --page raw --csv` (stdin or a file):
   python tools/ncu_summary.py raw.csv [nets per launch]
Columns: ms, DRAM GB read / written, DRAM % of peak, issue-slot %, FMA-pipe %, achieved warps %, registers, and thread-level
instructions per net (32 x warp instructions / nets, so inactive lanes count)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1] != "-" else sys.stdin))
nets = float(sys.argv[2]) if len(sys.argv) > 2 else None
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
names, units = rows[hdr], rows[hdr + 1]
col = {n: i for i, n in enumerate(names)}


def val(r, name, scale_units=None):
    if name not in col or r[col[name]] in ("", "n/a"):
        return float("nan")
    v = float(r[col[name]].replace(",", ""))
    u = units[col[name]].lower()
    if scale_units == "bytes":
        v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    if scale_units == "time":
        v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3, "nsecond": 1e-6}.get(u, 1)
    return v


print("%-62s %7s %7s %7s %6s %6s %6s %6s %4s %8s" % ("kernel", "ms", "GB rd", "GB wr", "dram%", "issue%", "fmaP%", "warps%", "regs",
                                                     "inst/net"))
for r in rows[hdr + 2:]:
    if len(r) < len(names):
        continue
    k = r[col["Kernel Name"]].replace("srnn::", "").replace("void ", "")
    k = k.split("(")[0][:62]
    inst = val(r, "smsp__inst_executed.sum")
    grid, block = val(r, "launch__grid_size"), val(r, "launch__block_size")
    per = 32.0 * inst / (nets if nets else grid * block)
    print("%-62s %7.3f %7.3f %7.3f %6.1f %6.1f %6.1f %6.1f %4d %8.0f" % (
        k, val(r, "gpu__time_duration.sum", "time"), val(r, "dram__bytes_read.sum", "bytes") / 1e9,
        val(r, "dram__bytes_write.sum", "bytes") / 1e9, val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        val(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), int(val(r, "launch__registers_per_thread")), per))
