import csv, io, subprocess, sys
keys=['gpu__time_duration.sum','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts.sum','l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed','smsp__warps_eligible.avg.per_cycle_active','l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed','l1tex__f_wavefronts.avg.pct_of_peak_sustained_elapsed','lts__t_sectors.avg.pct_of_peak_sustained_elapsed','l1tex__m_xbar2l1tex_read_sectors.avg.pct_of_peak_sustained_elapsed','l1tex__m_l1tex2xbar_write_sectors.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active','smsp__inst_executed_op_local_ld.sum','smsp__inst_executed_op_local_st.sum','sm__cycles_active.avg','smsp__cycles_active.avg']
for rep in sys.argv[1:]:
    raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
    rows=list(csv.reader(io.StringIO(raw))); hdr,units=rows[0],rows[1]
    for r in rows[2:]:
        print('==',rep,r[hdr.index('Kernel Name')][:60])
        for k in keys:
            if k in hdr: print('  ',k, r[hdr.index(k)], units[hdr.index(k)])
        st=[(h.split('stalled_')[1].split('_per')[0], float(r[i])) for i,h in enumerate(hdr) if 'warps_issue_stalled' in h and 'per_issue_active' in h and r[i] and float(r[i])>0.05]
        print('   stalls:', ' '.join(f'{a}={b:.2f}' for a,b in sorted(st,key=lambda x:-x[1])))
