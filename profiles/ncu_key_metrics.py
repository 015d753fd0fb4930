"""Print the handful of ncu raw-page metrics the profile notes quote.
usage: ncu_key_metrics.py report.ncu-rep [--index N (N-th captured launch, default 0)] [--traffic STAGE]"""
import csv, subprocess, sys

WANT = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed",
    "lts__t_sectors_srcunit_tex_op_read.sum",
    "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
IDX = int(sys.argv[sys.argv.index("--index") + 1]) if "--index" in sys.argv else 0
head, units, vals = rows[0], rows[1], rows[2 + IDX]
for n, u, v in zip(head, units, vals):
    if n == "Kernel Name": print(n, "=", v[:110])
    if n in WANT: print(f"{n} = {v} {u}")

# `--traffic STAGE`: record dram bytes read + written per launch of this kernel in profiles/ncu_traffic.json, the file
# bench.py's roofline.traffic is read from (stage names = bench.py's stage_ms keys)
if "--traffic" in sys.argv:
    import json, os
    stage = sys.argv[sys.argv.index("--traffic") + 1]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rec = dict(zip(head, zip(units, vals)))
    tot = sum(float(rec[m][1].replace(",", "")) * scale[rec[m][0]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json")
    try:
        d = json.load(open(path))
    except Exception:
        d = {}
    d[stage] = {"bytes_per_launch": tot, "kernel": rec["Kernel Name"][1][:60], "capture": os.path.basename(sys.argv[1]),
                "us": rec["gpu__time_duration.sum"][1]}
    json.dump(d, open(path, "w"), indent=1)
    print("wrote", path, d[stage])
