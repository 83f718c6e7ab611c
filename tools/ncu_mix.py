"""Dynamic instruction mix and stall totals of one kernel from `ncu --page source --csv --print-source sass`."""
import csv, collections, re, sys
rows=list(csv.reader(open(sys.argv[1])))
which=int(sys.argv[2]) if len(sys.argv)>2 else 1
hdr=None; data=[]
k=0
for r in rows:
    if r and r[0]=='Kernel Name':
        k+=1
        if k>which: break
        data=[]
        print(r[1][:100])
        continue
    if r and r[0]=='Address': hdr=r; continue
    if hdr and len(r)==len(hdr): data.append(r)
ix={h:i for i,h in enumerate(hdr)}
tot_s=sum(int(r[ix['# Samples']]) for r in data)
tot_i=sum(int(r[ix['Instructions Executed']]) for r in data)
print("static instrs",len(data),"samples",tot_s,"warp-instr executed",tot_i)
ops=collections.Counter(); sm=collections.Counter()
for r in data:
    m=re.match(r'\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)',r[ix['Source']])
    op=m.group(1) if m else '?'
    ops[op]+=int(r[ix['Instructions Executed']]); sm[op]+=int(r[ix['# Samples']])
print("dynamic mix (warp instr) top 25:")
for op,c in ops.most_common(25): print(f"  {op:12s} {c:12d} {100*c/tot_i:5.1f}%  samples {100*sm[op]/tot_s:5.1f}%")
for st in ['stall_wait','stall_no_inst','stall_short_sb','stall_branch_resolving','stall_math','stall_not_selected','stall_selected','stall_long_sb','stall_dispatch','stall_lg','stall_mio']:
    print(st, sum(int(r[ix[st]]) for r in data))
