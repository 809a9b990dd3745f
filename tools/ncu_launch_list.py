"""Aggregate an `ncu --csv` launch list of bench.py into the per-kernel table kept under profiles/.

Capture (on the GPU box, after the same command exited 0 without ncu):
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size \
        --clock-control none -s 592 -c 450 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline
Aggregate:
    python tools/ncu_launch_list.py gpurun_out/launches.csv > profiles/rNN_launches_step.md
The table covers the first complete train step inside the captured window (pack_input ... adam).
"""
import csv, sys, re, collections
rows=[]
with open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/launches.csv') as f:
    lines=[l for l in f if not l.startswith('==')]
rd=csv.DictReader(lines)
for r in rd: rows.append(r)
# group metric rows per launch ID
by=collections.OrderedDict()
for r in rows:
    k=r['ID']
    d=by.setdefault(k,{'name':r['Kernel Name']})
    d[r['Metric Name']]=(float(r['Metric Value'].replace(',','')), r['Metric Unit'])
ids=list(by)
names=[by[i]['name'] for i in ids]
# first complete step: from pack_input to adam
start=next(i for i,n in enumerate(names) if 'pack_input' in n)
end=next(i for i in range(start,len(names)) if 'adam' in names[i])
step=ids[start:end+1]
print('launches in step', len(step), file=sys.stderr)
def to_us(v,u): return v/1000 if u in('ns','nsecond') else (v if u in ('us','usecond') else v*1000)
def to_mb(v,u):
    m={'byte':1e-6,'Kbyte':1e-3,'Mbyte':1,'Gbyte':1e3}
    return v*m.get(u,1e-6)
agg=collections.OrderedDict()
tot=0
for i in step:
    d=by[i]
    n=re.sub(r'\(.*','',d['name'])
    n=n.replace('mrisr::','').replace('(anonymous namespace)::','').replace('tc::','')
    t=to_us(*d['gpu__time_duration.sum']); rdb=to_mb(*d['dram__bytes_read.sum']); wrb=to_mb(*d['dram__bytes_write.sum'])
    a=agg.setdefault(n,[0,0,0,0,d['launch__registers_per_thread'][0],0])
    a[0]+=1;a[1]+=t;a[2]+=rdb;a[3]+=wrb; a[5]+=d['sm__warps_active.avg.pct_of_peak_sustained_active'][0]*t
    tot+=t
print(f"| kernel | launches | us | share | DRAM read MB | DRAM write MB | GB/s | regs | occupancy % |")
print("|---|---|---|---|---|---|---|---|---|")
for n,a in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    if a[1]/tot<0.004: continue
    print(f"| {n} | {a[0]} | {a[1]:.0f} | {a[1]/tot:.3f} | {a[2]:.0f} | {a[3]:.0f} | {(a[2]+a[3])/a[1]*1e3:.0f} | {a[4]:.0f} | {a[5]/a[1]:.0f} |")
print(f"\ntotal {tot:.0f} us over {len(step)} launches")
