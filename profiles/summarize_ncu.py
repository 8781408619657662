"""Summarise .ncu-rep files (read with `ncu -i ... --page raw --csv`) into one text table:
per kernel duration, DRAM bytes, achieved DRAM GB/s, issue-slot and pipe utilisation, occupancy.
Usage: python profiles/summarize_ncu.py gpurun_out/r1_*.ncu-rep > profiles/r1_ncu_summary.txt"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_%"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu_%"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__inst_executed.sum", "warp_inst"),
]


def to_bytes(v, unit):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(v) * m.get(unit, 1)


def to_sec(v, unit):
    m = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1, "second": 1, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9}
    return float(v) * m.get(unit, 1)


for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        print(rep, "unreadable")
        continue
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")][:70]
        vals = {}
        for k, short in KEYS:
            if k in hdr:
                i = hdr.index(k)
                vals[short] = (r[i], units[i])
        t = to_sec(*vals["time"])
        rd, wr = to_bytes(*vals["dram_rd"]), to_bytes(*vals["dram_wr"])
        print(f"{rep.split('/')[-1]:28s} {name}")
        print(f"    time {t*1e3:9.3f} ms | dram read {rd/1e6:10.1f} MB write {wr/1e6:10.1f} MB -> {(rd+wr)/t/1e9:8.1f} GB/s "
              f"({vals['dram_%'][0][:5]}% of peak) | issue {vals['issue_%'][0][:5]}% alu {vals['alu_%'][0][:5]}% "
              f"fma {vals['fma_%'][0][:5]}% | warps active {vals['warps_%'][0][:5]}% | regs {vals['regs'][0]} "
              f"grid {vals['grid'][0]} x {vals['block'][0]} | warp-inst {float(vals['warp_inst'][0]):.3e}")
