O=gpurun_out; T=r2y
mkdir -p $O
F="ncu --set full --clock-control none --import-source on -f"
timeout 600 $F -k regex:msm_accumulate -s 3 -c 1 -o $O/${T}_acc python bench.py --steps 2 --warmup 3 --no-extras > $O/${T}_ncu_acc.log 2>&1; echo "ncu acc $?"
python tools/ncu_summary.py $O/${T}_acc.ncu-rep > $O/${T}_acc_summary.txt 2>&1
ncu -i $O/${T}_acc.ncu-rep --page source --csv > $O/${T}_acc_source.csv 2>/dev/null; python - <<'PY'
import csv,re,collections
rows=list(csv.reader(open('gpurun_out/r2y_acc_source.csv')))
segs=[];cur=None
for r in rows:
    if r and r[0]=='Kernel Name': cur=[];segs.append(cur);continue
    if cur is not None: cur.append(r)
seg=segs[0]; hdr=seg[0]; ix={h:i for i,h in enumerate(hdr)}
stall=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot=collections.Counter(); ex=collections.Counter()
for r in seg[1:]:
    if len(r)<len(hdr): continue
    src=r[ix['Source']].strip(); op=re.sub(r'^@!?U?P\d+\s+','',src).split()[0] if src else '?'
    k=op.split('.')[0]+('.WIDE' if 'WIDE' in op else '')+('.HI' if '.HI' in op else '')
    try: e=int(r[ix['Instructions Executed']] or 0)
    except: continue
    ex[k]+=e
    for c in stall: tot[c]+=int(r[ix[c]] or 0)
S=sum(tot.values()); E=sum(ex.values())
with open('gpurun_out/r2y_acc_mix.txt','w') as f:
    f.write(f"msm_accumulate_kernel<Fq>: warp-stall samples {S}, warp instructions executed {E}\n")
    for c,v in tot.most_common(8): f.write(f"  {c:24s} {v/S:.3f}\n")
    for k,v in ex.most_common(10): f.write(f"  {k:14s} {v/E:.3f} of executed instructions\n")
print(open('gpurun_out/r2y_acc_mix.txt').read())
PY
rm -f $O/${T}_acc.ncu-rep $O/${T}_acc_source.csv
