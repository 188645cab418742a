"""Summarise an .ncu-rep (ncu --set full) into a small JSON: per kernel launch, the metrics DESIGN.md / bench.py quote.
Usage: python tools/ncu_summary.py report.ncu-rep > profiles/summary.json"""
import csv, io, json, subprocess, sys

WANT = {
    "gpu__time_duration.sum": "time_us",
    "dram__bytes_read.sum": "dram_read_bytes",
    "dram__bytes_write.sum": "dram_write_bytes",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
    "lts__t_bytes.sum": "l2_bytes",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct_of_peak",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct_of_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": "tensor_pipe_pct_of_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct_of_peak",
    "sm__cycles_elapsed.max": "sm_cycles",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "launch__registers_per_thread": "registers",
    "launch__occupancy_limit_registers": "occ_limit_regs",
    "launch__grid_size": "grid",
    "launch__cluster_size": "cluster",
}
UNIT_SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1, "us": 1, "ms": 1e3, "ns": 1e-3, "second": 1e6}


def main():
    raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = []
    for r in data:
        rec = {"kernel": r[hdr.index("Kernel Name")].split("(")[0].split("::")[-1]}
        for i, h in enumerate(hdr):
            if h in WANT and r[i] != "":
                try:
                    v = float(r[i].replace(",", ""))
                except ValueError:
                    continue
                v *= UNIT_SCALE.get(units[i], 1)
                rec[WANT[h]] = v
        if "dram_read_bytes" in rec and "dram_write_bytes" in rec:
            rec["dram_bytes"] = rec["dram_read_bytes"] + rec["dram_write_bytes"]
            if rec.get("time_us"):
                rec["dram_GBps"] = round(rec["dram_bytes"] / rec["time_us"] / 1e3, 1)
        out.append(rec)
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
