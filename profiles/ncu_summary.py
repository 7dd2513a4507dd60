"""Per-kernel summary of an `ncu --set full` capture, as committed in profiles/r01_ncu_full_*.json.

Capture (one GPU, after the same command has exited 0 without ncu):
  FDLS_BENCH_FRAMES_PER_STEP=2 ncu --set full --clock-control none --import-source on \
      -k regex:"preprocess_fwd|rs_onesweep|scan_|part_|render_fwd" -s 22 -c 11 -o gpurun_out/prof \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline
Summarise (here):
  ncu -i gpurun_out/prof.ncu-rep --page raw --csv > /tmp/raw.csv && python profiles/ncu_summary.py /tmp/raw.csv out.json
"""
import csv
import json
import sys

WANT = {"t": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
        "sm": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l2hit": "lts__t_sector_hit_rate.pct",
        "occ": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
        "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "inst": "smsp__inst_executed.sum",
        "grid": "launch__grid_size"}
SCALE = {"Kbyte": 1e-3, "byte": 1e-6, "Mbyte": 1.0, "Gbyte": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3}  # -> MB, us


def main(raw_csv, out_json):
    rows = list(csv.reader(open(raw_csv)))
    hdr, units = rows[0], rows[1]
    out = []
    print("%-28s %7s %7s %7s %5s %6s %6s %5s %4s %6s %7s %6s" % ("kernel", "us", "rdMB", "wrMB", "SM%", "DRAM%", "L2hit",
                                                                 "occ%", "regs", "issue%", "Minst", "grid"))
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").replace("fdls::", "")
        d = {}
        for k, m in WANT.items():
            i = hdr.index(m)
            d[k] = float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)
        out.append({"kernel": name, **d})
        print("%-28s %7.1f %7.2f %7.2f %5.1f %6.1f %6.1f %5.1f %4.0f %6.1f %7.2f %6.0f" % (
            name[:28], d["t"], d["rd"], d["wr"], d["sm"], d["dram"], d["l2hit"], d["occ"], d["regs"], d["issue"],
            d["inst"] / 1e6, d["grid"]))
    json.dump(out, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
