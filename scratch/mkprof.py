import csv, re, gzip, json, subprocess, sys
tag = sys.argv[1]  # e.g. p3
out = sys.argv[2]  # e.g. r1b
lines=[l for l in open(f'gpurun_out/{tag}_launches.csv') if l.startswith('"')]
rows=list(csv.DictReader(lines))
def short(n):
    m=re.match(r'(?:void )?(?:dli::)?([a-z_0-9]+)(<[^(]*>)?', n); return (m.group(1)+(m.group(2) or '')) if m else n[:40]
with open(f'profiles/{out}_launches_bench_steps4.csv','w') as f:
    f.write('# ncu --metrics gpu__time_duration.sum --clock-control none -k <our kernels>; python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extra (B200)\n')
    f.write('id,kernel,grid,block,duration_ns\n')
    for x in rows:
        f.write(f"{x['ID']},{short(x['Kernel Name'])},{x['Grid Size'].replace(', ','x')},{x['Block Size'].replace(', ','x')},{x['Metric Value']}\n")
idx=[i for i,x in enumerate(rows) if x['Kernel Name'].startswith('embedding_kernel') and x['Grid Size']=='(1, 1, 1)']
pidx=[i for i,x in enumerate(rows) if x['Kernel Name'].startswith('embedding_kernel') and x['Grid Size']=='(2048, 1, 1)']
o=[]
def summarize(seg,title):
    agg={}
    for x in seg:
        k=short(x['Kernel Name'])+' grid'+x['Grid Size']
        a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=float(x['Metric Value'])/1e3
    tot=sum(a[1] for a in agg.values())
    o.append(f'{title}: {len(seg)} launches, {tot:.1f} us summed kernel time (ncu: cold-cache, serialised; compare SHARES)')
    for k,(n,t) in sorted(agg.items(), key=lambda kv:-kv[1][1]):
        o.append('  %8.1f us %5.1f%%  n=%3d  avg %7.2f us  %s'%(t,100*t/tot,n,t/n,k))
summarize(rows[idx[-2]:idx[-1]],'one decode step (Llama-3-8B, B=1, ctx 2048)')
i0=pidx[-1]; i1=[i for i in idx if i>i0][0]
summarize(rows[i0:i1],'one prefill (Llama-3-8B, B=1, T=2048)')
open(f'profiles/{out}_launch_shares.txt','w').write('\n'.join(o)+'\n')
print('\n'.join(o))
want=['Kernel Name','Grid Size','Block Size','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_tensor_op_hmma.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','sm__throughput.avg.pct_of_peak_sustained_elapsed','launch__shared_mem_per_block_dynamic','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct']
with open(f'profiles/{out}_ncu_full_raw.txt','w') as f:
    f.write('# ncu --set full --clock-control none --import-source on (B200). Raw-page excerpts per captured launch.\n')
    for k in ('gemv','gemm','attn'):
        try:
            txt=subprocess.run(['ncu','-i',f'gpurun_out/{tag}_{k}.ncu-rep','--page','raw','--csv'],capture_output=True,text=True).stdout
        except Exception as e:
            continue
        r=list(csv.reader(txt.splitlines()))
        if len(r)<3: continue
        hdr,units=r[0],r[1]
        ii=[(w,hdr.index(w)) for w in want if w in hdr]
        f.write(f'== {k}\n')
        for row in r[2:]:
            f.write('---\n')
            for w,i in ii: f.write(f'{w} [{units[i]}]: {row[i][:150]}\n')
subprocess.run(f'gzip -c gpurun_out/{tag}_launches.csv > profiles/{out}_launches_raw.csv.gz', shell=True)
