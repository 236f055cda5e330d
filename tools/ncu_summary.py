"""Prints the key metrics of every kernel in an .ncu-rep (run here, no GPU): python tools/ncu_summary.py file.ncu-rep"""
import csv
import io
import subprocess
import sys

WANT = {
    "gpu__time_duration.sum": "time",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor%",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "lts%",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": "l1%",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm%",
    "dram__bytes_read.sum": "dramR",
    "dram__bytes_write.sum": "dramW",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram%",
    "l1tex__m_xbar2l1tex_read_bytes.sum": "l2->sm",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": "memtensor%",
    "smsp__cycles_active.avg": "cyc_act",
    "launch__registers_per_thread": "regs",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "lsu_smem%",
    "smsp__inst_executed.sum": "inst",
}


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        d = {}
        for i, h in enumerate(hdr):
            if h in WANT:
                v = r[i]
                try:
                    v = f"{float(v.replace(',', '')):.4g}"
                except ValueError:
                    pass
                d[WANT[h]] = f"{v}{units[i] if units[i] not in ('%',) else ''}"
        print(name[:60])
        print("   ", "  ".join(f"{k}={v}" for k, v in d.items()))


if __name__ == "__main__":
    main(sys.argv[1])
