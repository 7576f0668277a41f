"""Key numbers of one ncu report: python tools/ncu_brief.py file.ncu-rep [sections_per_launch [launch index]]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
row = int(sys.argv[3]) if len(sys.argv) > 3 else 0
d = dict(zip(rows[0], rows[2 + row]))
print('kernel', d.get('Kernel Name'), 'grid', d.get('Grid Size'), 'block', d.get('Block Size'))
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
fp32 = g("smsp__sass_thread_inst_executed_op_fp32_pred_on.sum")
print(f"fp32 thread-instr (scalar fadd+fmul+ffma) {fp:.4g}; all FP32 incl. packed FADD2/FMUL2/FFMA2 {fp32:.4g} = {fp32/32:.4g} warp-instr "
      f"({fp32/32/inst*100:.0f}% of all)")
print(f"pipes: fma {g('sm__inst_executed_pipe_fma.sum'):.4g} warp-instr, fma cycles active {g('sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active'):.1f}% "
      f"(heavy {g('sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active'):.1f}%), alu {g('sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active'):.1f}%, "
      f"lsu {g('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active'):.1f}%, issue active {g('sm__issue_active.avg.pct_of_peak_sustained_elapsed'):.1f}%, "
      f"eligible warps/cycle {g('smsp__warps_eligible.avg.per_cycle_active'):.2f}")
st = {k[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]: float(v) for k, v in d.items()
      if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio')}
print("stalls/issue: " + ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:9]))
if len(sys.argv) > 2:
    n = float(sys.argv[2])
    print(f"per section: {inst/n:.0f} warp-instr, {cyc*148/n:.0f} SM-cycles, {g('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum')/n:.0f} smem wavefronts")
