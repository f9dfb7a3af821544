#!/usr/bin/env python3
"""Summarise an ncu report: key raw metrics + SASS opcode mix + stall reasons (dev helper)."""
import csv, collections, re, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(raw)))
hdr,units=rows[0],rows[1]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','smsp__inst_executed.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','sm__inst_executed_pipe_fma.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__cycles_elapsed.max','smsp__issue_active.avg.pct_of_peak_sustained_active','lts__t_sectors_srcunit_tex_op_read.sum','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','launch__grid_size','launch__waves_per_multiprocessor']
for r in rows[2:]:
    d=dict(zip(hdr,r))
    print("==",d.get('Kernel Name','')[:60])
    for k in hdr:
        if any(k==w or (k.startswith(w) and k[len(w):] in ('', '.pct_of_peak_sustained_elapsed','.pct_of_peak_sustained_active')) for w in want):
            print(f"  {k:75s} {d[k]:>16s} {units[hdr.index(k)]}")
src = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(src)))
hdr=rows[1]
iA=hdr.index('Source'); iN=hdr.index('Instructions Executed')
ops=collections.Counter(); tot=0; n=0
stall_cols=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
st=collections.Counter()
for r in rows[2:]:
    if len(r)<len(hdr):
        if r and r[0]=='Kernel Name': break
        continue
    try: cnt=int(r[iN])
    except: continue
    m=re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)',r[iA]); op=m.group(2).split('.')[0] if m else '?'
    ops[op]+=cnt; tot+=cnt; n+=1
    for i in stall_cols:
        try: st[hdr[i]]+=int(r[i])
        except: pass
print('static instrs',n,'dynamic',tot)
print('  '.join(f"{op}:{c/tot*100:.1f}%" for op,c in ops.most_common(16)))
ts=sum(st.values())
print('  '.join(f"{k[6:]}:{v/ts*100:.0f}%" for k,v in st.most_common(10)))
