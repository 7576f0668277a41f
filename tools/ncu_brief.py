"""Key numbers of one ncu report: python tools/ncu_brief.py file.ncu-rep [sections_per_launch]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = dict(zip(rows[0], rows[2]))
def g(k):
    try: return float(d[k].replace(",", ""))
    except Exception: return float("nan")
cyc = g("sm__cycles_elapsed.max")
inst = g("smsp__inst_executed.sum")
print(f"duration {g('gpu__time_duration.sum'):.4g} {rows[1][rows[0].index('gpu__time_duration.sum')]}, cycles {cyc:.0f}, warp-instr {inst:.3g}, regs {d.get('launch__registers_per_thread')}, "
      f"issue busy {g('sm__inst_issued.avg.pct_of_peak_sustained_active') if 'sm__inst_issued.avg.pct_of_peak_sustained_active' in d else float('nan'):.1f}%, ipc {g('sm__inst_executed.avg.per_cycle_active'):.2f}")
print(f"smem wavefronts {g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum'):.3g} ({g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed'):.1f}% of peak), "
      f"bank conflicts ld {g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum'):.3g} st {g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum'):.3g}, "
      f"shared instr {g('smsp__sass_inst_executed_op_shared.sum'):.3g}, local ld {g('smsp__sass_inst_executed_op_local_ld.sum'):.3g}")
print(f"dram read {g('dram__bytes_read.sum')} {rows[1][rows[0].index('dram__bytes_read.sum')]}, write {g('dram__bytes_write.sum')} {rows[1][rows[0].index('dram__bytes_write.sum')]}")
fp = sum(g(f"smsp__sass_thread_inst_executed_op_{o}_pred_on.sum") for o in ("fadd", "fmul", "ffma"))
print(f"fp32 thread-instr {fp:.4g} = {fp/32:.4g} warp-instr ({fp/32/inst*100:.0f}% of all)")
st = {k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]: float(v) for k, v in d.items()
      if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio')}
print("stalls/issue: " + ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:9]))
if len(sys.argv) > 2:
    n = float(sys.argv[2])
    print(f"per section: {inst/n:.0f} warp-instr, {cyc*148/n:.0f} SM-cycles, {g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum')/n:.0f} smem wavefronts")
