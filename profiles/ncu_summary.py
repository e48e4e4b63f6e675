import csv,sys
from collections import Counter
rep=sys.argv[1]
import subprocess
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
for vals in rows[2:]:
    d={h:(v,u) for h,u,v in zip(hdr,units,vals)}
    print('==', d['Kernel Name'][0][:80])
    keys=['gpu__time_duration.sum','launch__registers_per_thread','launch__grid_size','launch__block_size','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__warps_eligible.avg.per_cycle_active','smsp__warps_active.avg.per_cycle_active','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','dram__bytes_read.sum','dram__bytes_write.sum','smsp__inst_executed_op_local_ld.sum','smsp__inst_executed_op_local_st.sum']
    for k in keys: print(' ',k, d.get(k))
    st=[(float(v[0].replace(',','')),k) for k,v in d.items() if 'smsp__average_warps_issue_stalled' in k and k.endswith('_per_issue_active.ratio')]
    for v,k in sorted(st,reverse=True)[:8]: print('  stall',round(v,2),k.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''))
sass=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','sass'],capture_output=True,text=True).stdout
rows=list(csv.reader(sass.splitlines()))
hdr=rows[1]; data=[r for r in rows[2:] if len(r)==len(hdr)]
iE=hdr.index('Instructions Executed'); iS=hdr.index('Source'); iA=hdr.index('Address')
ex=[(int(r[iA],16), r[iS], int(r[iE])) for r in data]
tot=sum(e[2] for e in ex); hot=[e for e in ex if e[2]>0]
print('instr total',len(ex),'executed',len(hot),'=',len(hot)*16/1024,'KB')
h=sorted(hot,key=lambda e:-e[2]); acc=0
for frac in (0.9,0.99):
    acc=0
    for i,e in enumerate(h):
        acc+=e[2]
        if acc>frac*tot: print(frac,'of executions in',i+1,'instr =',(i+1)*16/1024,'KB'); break
c=Counter()
for e in hot:
    t=e[1].split()
    op=t[1] if t[0].startswith('@') else t[0]
    c[op.split('.')[0]]+=e[2]
print(' '.join('%s:%.1f%%'%(k,100*v/tot) for k,v in c.most_common(18)))
