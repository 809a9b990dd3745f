"""Aggregate an `ncu --csv` launch list of bench.py into the per-kernel table kept under profiles/, and (--json OUT) the per-family
DRAM traffic record bench.py reads for roofline.traffic (tied to the source tree by bench.source_hash()).

Capture (on the GPU box, after the same command exited 0 without ncu):
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size \
        --clock-control none -s 592 -c 450 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline
Aggregate:
    python tools/ncu_launch_list.py gpurun_out/launches.csv > profiles/rNN_launches_step.md
The table covers the first complete train step inside the captured window (pack_input ... adam).
"""
import csv, sys, re, collections, json, os
args=[a for a in sys.argv[1:] if not a.startswith('--')]
json_out=None
if '--json' in sys.argv:
    json_out=sys.argv[sys.argv.index('--json')+1]
    args=[a for a in args if a!=json_out]
rows=[]
with open(args[0] if args else 'gpurun_out/launches.csv') as f:
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
# last complete step inside the window: from pack_input to the Adam update
starts=[i for i,n in enumerate(names) if 'pack_input' in n]
step=None
for st in reversed(starts):
    ends=[i for i in range(st,len(names)) if 'adam_dev_kernel' in names[i] or 'adam_kernel' in names[i]]
    if ends:
        step=ids[st:ends[0]+1]; break
assert step, 'no complete train step in the capture' 
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

# kernel-name -> bench.py family (the profiler tags of csrc/*.cu)
FAMILIES=[('gn_bwd',r'gn_bwd_'),('gn_apply',r'gn_apply_kernel'),('l1_ssim',r'l1_ssim_kernel|ssim_fwd_maps|ssim_grad_kernel|ssim_rows|ssim_cols|loss_finalize|ssim_'),
          ('upsample2x_bwd',r'upsample2x_bwd'),('upsample2x',r'upsample2x'),('conv_ring',r'conv_ring_kernel'),('conv_tc',r'conv_tc_kernel'),
          ('wgrad_rows',r'wgrad_rows'),('wgrad_tc3',r'wgrad_tc3'),('wgrad_tc',r'wgrad_tc_kernel|wgrad_tc_reduce'),('wgrad_ring',r'wgrad_ring'),
          ('conv_first',r'conv_first_kernel'),('wgrad_first',r'wgrad_first'),('head_blend',r'head_blend'),('final_head',r'final_head'),
          ('colsum',r'colsum'),('pack_weights',r'pack_w'),('gn_stats',r'gn_stats'),('gn_finalize',r'gn_finalize'),('adam',r'adam')]
if json_out:
    fam=collections.OrderedDict()
    for n,a in agg.items():
        for f,rx in FAMILIES:
            if re.search(rx,n):
                fam[f]=fam.get(f,0.0)+(a[2]+a[3])*1e6; break
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    rec={'source_hash':bench.source_hash(),'workload':os.environ.get('MRISR_TRAFFIC_WORKLOAD','train:64x256x256'),
         'what':'dram__bytes_read.sum + dram__bytes_write.sum per train step, summed over the launches of each kernel family (ncu --clock-control none, eager launches)',
         'capture':os.path.basename(args[0] if args else 'launches.csv'),'step_us_under_ncu':tot,'launches':len(step),'families':fam}
    json.dump(rec,open(json_out,'w'),indent=1)
