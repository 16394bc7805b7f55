"""Extract the metrics DESIGN.md / bench.py cite from an .ncu-rep (needs `ncu` on PATH) -> CSV on stdout."""
import csv, re, subprocess, sys
KEEP = r"^(gpu__time_duration\.sum|sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|dram__bytes_(read|write)\.sum(\.per_second)?|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|lts__t_sector_hit_rate\.pct|l1tex__m_xbar2l1tex_read_bytes\.sum(\.per_second)?|lts__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__cycles_active\.avg|sm__cycles_elapsed\.max|launch__registers_per_thread|launch__grid_size|launch__block_size|smsp__inst_executed\.sum|sm__warps_active\.avg\.pct_of_peak_sustained_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|launch__shared_mem_per_block_dynamic)$"
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
name_i = hdr.index("Kernel Name")
cols = [i for i, h in enumerate(hdr) if re.match(KEEP, h)]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [re.sub(r"^.*?(Epi\w+|bwd_persistent_kernel|l2norm_rows_kernel|[a-z_]+_kernel).*$", r"\1", r[name_i]) for r in data])
for i in cols:
    w.writerow([hdr[i], units[i]] + [r[i] for r in data])
