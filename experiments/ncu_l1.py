"""L1 / shared data-pipe breakdown of every kernel in an ncu report (wavefronts per kind, % of peak, conflicts).
usage: python experiments/ncu_l1.py report.ncu-rep [units]   (units: divide the wavefront counts, e.g. frames)"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
want = [
    ("lsu wavefronts % peak", "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed"),
    ("shared ld", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum"),
    ("shared st", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum"),
    ("shared conflicts ld", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum"),
    ("shared conflicts st", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum"),
    ("global ld wavefronts", "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum"),
    ("global st wavefronts", "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_st.sum"),
    ("global ld sectors", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"),
    ("global st sectors", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum"),
    ("global ld requests", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"),
    ("global st requests", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum"),
    ("lsu writeback active %", "l1tex__lsu_writeback_active.sum.pct_of_peak_sustained_elapsed"),
    ("inst executed", "smsp__inst_executed.sum"),
    ("cycles", "sm__cycles_elapsed.max"),
    ("dram read B", "dram__bytes_read.sum"),
    ("dram write B", "dram__bytes_write.sum"),
]
for r in rows[2:]:
    print(r[col["Kernel Name"]][:60])
    for name, key in want:
        if key in col and r[col[key]]:
            v = float(r[col[key]].replace(",", ""))
            print(f"   {name:26s} {v:16.1f}" + (f"   per unit {v / units:10.2f}" if units != 1 and "%" not in name else ""))
