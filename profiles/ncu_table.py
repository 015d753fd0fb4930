"""Markdown table of every kernel in an ncu report (one `--set full` capture of a whole simulation wave).
usage: ncu_table.py report.ncu-rep"""
import csv
import subprocess
import sys

COLS = [
    ("gpu__time_duration.sum", "us", lambda v, u: v / 1e3 if u in ("ns", "nsecond") else v),
    ("dram__bytes_read.sum", "MB read", None),
    ("dram__bytes_write.sum", "MB written", None),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak", None),
    ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue slots %", None),
    ("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %", None),
    ("sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "UMMA bf16 % of peak", None),
    ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem operand wavefronts %", None),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %", None),
    ("launch__registers_per_thread", "regs", None),
    ("launch__grid_size", "grid", None),
    ("launch__block_size", "block", None),
]


def to_bytes(v, u):
    u = u.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units, body = rows[0], rows[1], rows[2:]
    ix = {n: i for i, n in enumerate(head)}
    seen = {}
    for r in body:
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("gz::", "")
        rec = seen.setdefault(name, {"n": 0, "vals": {}})
        rec["n"] += 1
        for m, _, _ in COLS:
            if m not in ix or r[ix[m]] in ("", "n/a"):
                continue
            v, u = float(r[ix[m]].replace(",", "")), units[ix[m]]
            if m.startswith("dram__bytes"):
                v = to_bytes(v, u) / 1e6
            elif m == "gpu__time_duration.sum":
                v = v / 1e3 if u.startswith("n") else v * 1e3 if u.startswith("m") else v
            rec["vals"].setdefault(m, []).append(v)
    print("| kernel | captured | " + " | ".join(c[1] for c in COLS) + " | DRAM GB/s |")
    print("|---|---|" + "---|" * (len(COLS) + 1))
    for name, rec in seen.items():
        cells = []
        avg = {}
        for m, _, _ in COLS:
            vs = rec["vals"].get(m)
            avg[m] = sum(vs) / len(vs) if vs else None
            cells.append("%.1f" % avg[m] if vs else "-")
        t = avg.get("gpu__time_duration.sum")
        by = (avg.get("dram__bytes_read.sum") or 0) + (avg.get("dram__bytes_write.sum") or 0)
        cells.append("%.0f" % (by * 1e6 / (t * 1e-6) / 1e9) if t else "-")
        print("| %s | %d | " % (name[:60], rec["n"]) + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
