"""Condense `ncu -i X.ncu-rep --page raw --csv` into the handful of metrics DESIGN.md / bench.py cite."""
import csv, sys
WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (TriageCompute)"),
    ("sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", "bf16 tensor ops % of peak"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor-memory (TMEM) cycles active %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "LSU shared-memory wavefronts"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1/shared throughput %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("sm__cycles_elapsed.avg", "SM cycles elapsed"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__inst_executed.sum", "warp instructions"),
]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {}
for i, h in enumerate(hdr):
    idx.setdefault(h.split(".TriageCompute.")[-1] if "TriageCompute" in h else h, i)
for r in rows[2:]:
    print("## %s" % r[idx["Kernel Name"]].split("(")[0])
    for key, label in WANT:
        if key in idx:
            print("  %-44s %s %s" % (label, r[idx[key]], units[idx[key]]))
    print()
