"""Compact per-launch table from an ncu report: `python tools/ncu_summary.py rep.ncu-rep [regex]`.
Columns: duration, tensor-pipe active %, SM clock, DRAM read/write bytes, DRAM throughput %, registers, smem."""
import csv, io, re, subprocess, sys

COLS = [("gpu__time_duration.sum", "dur"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%act"),
        ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor%elapsed"),
        ("sm__cycles_elapsed.avg.per_second", "smclk"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"), ("lts__t_sector_hit_rate.pct", "l2hit%")]


def main():
    rep = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, rows = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    idx = [(hdr.index(m), n) for m, n in COLS if m in hdr]
    print("| kernel | " + " | ".join(n for _, n in idx) + " |")
    print("|---|" + "---|" * len(idx))
    for r in rows:
        name = re.sub(r"^void |<unnamed>::|\(.*$", "", r[ki])
        if pat and not pat.search(r[ki]):
            continue
        print("| %s | " % name[:48] + " | ".join("%s %s" % (r[i], units[i]) for i, _ in idx) + " |")


if __name__ == "__main__":
    main()
