"""Per-launch summary of an `ncu --set full` capture: reads the `ncu -i rep --page raw --csv` dump and writes
kernel,grid,duration_us,dram_bytes_read,dram_bytes_write,tensor_pipe_active_pct,registers (bytes in bytes, whatever
unit ncu chose for the column).  Usage: python tools/ncu_full_summary.py raw.csv > profiles/rNN_ncu_gemm_full.csv"""
import csv
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0,
         "nsecond": 1e-3, "msecond": 1e3}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[hi], rows[hi + 1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, name, kind):
        i = ix[name]
        v = float(r[i].replace(",", ""))
        return v * SCALE.get(units[i], 1.0) if kind else v

    print("kernel,grid,duration_us,dram_bytes_read,dram_bytes_write,tensor_pipe_active_pct,registers")
    for r in rows[hi + 2:]:
        if len(r) != len(hdr):
            continue
        tp = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"
        print('"%s",%s,%.2f,%d,%d,%s,%s' % (
            r[ix["Kernel Name"]].split("(")[0], r[ix["launch__grid_size"]], val(r, "gpu__time_duration.sum", True),
            val(r, "dram__bytes_read.sum", True), val(r, "dram__bytes_write.sum", True),
            r[ix[tp]] if tp in ix else "", r[ix["launch__registers_per_thread"]]))


if __name__ == "__main__":
    main()
