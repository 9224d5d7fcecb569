"""Markdown table of the key counters of every kernel in an `ncu --set full` report:
   python tools/ncu_full_table.py <report.ncu-rep>   (reads it with `ncu -i ... --page raw --csv`)"""
import csv
import io
import subprocess
import sys

COLS = [("us", "gpu__time_duration.sum", 1e-3, "%.1f"),
        ("warp inst (M)", "smsp__inst_executed.sum", 1e-6, "%.1f"),
        ("issue active %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1, "%.1f"),
        ("tensor pipe %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1, "%.1f"),
        ("dram rd MB", "dram__bytes_read.sum", None, "%.1f"),
        ("dram wr MB", "dram__bytes_write.sum", None, "%.1f"),
        ("dram % of peak", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1, "%.1f"),
        ("L2->SM GB", "l1tex__m_xbar2l1tex_read_bytes.sum", None, "%.2f"),
        ("tma ld / st inst", None, None, None),
        ("regs", "launch__registers_per_thread", 1, "%d"),
        ("smem KB", "launch__shared_mem_per_block_dynamic", None, "%.0f"),
        ("grid x block", None, None, None)]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1.0, "us": 1e3, "ms": 1e6, "Kbyte/block": 1e3, "byte/block": 1.0, "s": 1e9,
        "usecond": 1e3, "nsecond": 1.0, "msecond": 1e6, "second": 1e9}


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, name, want=None):
        if name not in idx:
            return None
        v = r[idx[name]].replace(",", "")
        if v in ("", "n/a"):
            return None
        f = float(v)
        u = units[idx[name]]
        if u in UNIT:
            f *= UNIT[u]
        return f

    print("| kernel | " + " | ".join(c[0] for c in COLS) + " |")
    print("|---|" + "---:|" * len(COLS))
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("dmx::", "")
        cells = []
        for title, metric, scale, fmt in COLS:
            if title == "tma ld / st inst":
                a, b = val(r, "smsp__inst_executed_op_tma_ld.sum"), val(r, "smsp__inst_executed_op_tma_st.sum")
                cells.append("%d / %d" % (a or 0, b or 0))
            elif title == "grid x block":
                cells.append("%s x %s" % (r[idx["Grid Size"]].replace(" ", ""), r[idx["Block Size"]].replace(" ", "")))
            else:
                v = val(r, metric)
                if v is None:
                    cells.append("-")
                elif title.endswith("MB"):
                    cells.append(fmt % (v / 1e6))
                elif title.endswith("GB"):
                    cells.append(fmt % (v / 1e9))
                elif title == "smem KB":
                    cells.append(fmt % (v / 1e3))
                else:
                    cells.append(fmt % (v * scale))
        print("| `%s` | " % name + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
