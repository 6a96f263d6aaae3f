import csv, subprocess, sys
rep=sys.argv[1]
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr=rows[0]
keys=['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__t_sector_hit_rate.pct','sm__throughput.avg.pct_of_peak_sustained_elapsed','sm__inst_executed.avg.per_cycle_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_tensor.sum','smsp__inst_executed.sum']
for r in rows[2:]:
    d=dict(zip(hdr,r))
    print('----')
    for k in keys:
        if k in d: print(k, d[k][:100])
    for k,v in d.items():
        if 'issue_stalled' in k and 'per_issue_active' in k:
            try:
                if float(v)>0.3: print('  stall', k.split('issue_stalled_')[1].split('_per')[0], v)
            except: pass
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
his=[i for i,r in enumerate(rows) if r and r[0]=='Address']
for hi in his[:int(sys.argv[2]) if len(sys.argv)>2 else 1]:
    hdr=rows[hi]; body=[]
    for r in rows[hi+1:]:
        if r and r[0]=='Kernel Name': break
        body.append(r)
    si=hdr.index('Warp Stall Sampling (All Samples)'); ii=hdr.index('Instructions Executed')
    tot=sum(int(r[si]) for r in body)
    print('total samples',tot,'n instr',len(body))
    for k,r in enumerate(body):
        s=int(r[si])
        if s>tot*0.008: print(k, r[1].strip()[:80], s, f'{100*s/tot:.1f}%', r[ii])
