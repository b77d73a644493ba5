#!/usr/bin/env python
"""Per-source-line stall samples and instruction counts of one kernel from an .ncu-rep (ncu --set full
--import-source on): joins ncu's SASS page (by instruction address) with the line table nvdisasm -g prints
for the cubin of the same build.

    cuobjdump -xelf all operator_b200/csrc/pm_kernels.o          # -> pm_kernels.sm_100a.cubin (run in /tmp/cub)
    python tools_ncu_hotspots.py k_planE gpurun_out/post_r02.ncu-rep 30 [ncu kernel-name regex, default: the mangled fragment]

The object file must be the build that was profiled (line numbers refer to its source)."""
import csv,re,subprocess,sys,collections
kern=sys.argv[1]   # e.g. k_planE
rep=sys.argv[2]
dis=subprocess.run("nvdisasm -g -c /tmp/cub/pm_kernels.sm_100a.cubin",shell=True,capture_output=True,text=True).stdout.splitlines()
# locate function
start=None
for i,l in enumerate(dis):
    if l.startswith('.text.') and kern in l and l.endswith(':'): start=i
lines=[]; cur=None; inl=None
for l in dis[start+1:]:
    if l.startswith('//--------------------- .text.') or l.startswith('//--------------------- .'):
        if lines: break
    m=re.search(r'//## File "(.*?)", line (\d+)(.*)',l)
    if m:
        f=m.group(1); cur=int(m.group(2)); extra=m.group(3)
        m2=re.search(r'inlined at "(.*?)", line (\d+)',extra)
        inl=int(m2.group(2)) if m2 else None
        if not f.endswith('pm_kernels.cu'):
            cur = inl if (m2 and m2.group(1).endswith('pm_kernels.cu')) else -1
        continue
    m=re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);',l)
    if m: lines.append((int(m.group(1),16),cur,inl,m.group(2)))
kre=sys.argv[4] if len(sys.argv)>4 else kern[:-1]
out=subprocess.run(f"ncu -i {rep} --page source --csv --kernel-name regex:{kre}",shell=True,capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
for i,r in enumerate(rows):
    if '# Samples' in r: hdr=r; st=i+1; break
iN=hdr.index('# Samples'); iI=hdr.index('Instructions Executed')
data=rows[st:]
base=int(data[0][0],16)
src=open('/root/repo/operator_b200/csrc/pm_kernels.cu').read().splitlines()
agg=collections.defaultdict(lambda:[0,0])
tot=[0,0]
bymap={off:(ln,inl,txt) for off,ln,inl,txt in lines}
seen=set()
for r in data:
    if not r or not r[0].startswith('0x'): continue
    off=int(r[0],16)-base
    if off in seen or off not in bymap: continue
    seen.add(off)
    ln,inl,txt=bymap[off]
    n=int(r[iN]); ie=int(r[iI])
    agg[ln][0]+=n; agg[ln][1]+=ie; tot[0]+=n; tot[1]+=ie
print('total samples',tot[0],'instr',tot[1],'n sass',len(data),len(lines))
for ln,(n,ie) in sorted(agg.items(),key=lambda kv:-kv[1][0])[:int(sys.argv[3]) if len(sys.argv)>3 else 30]:
    print(f"{n:6d} {100*n/tot[0]:5.1f}%  inst {ie:9d} {100*ie/tot[1]:5.1f}%  L{ln}: {src[ln-1].strip()[:110] if ln and ln>0 else '(header)'}")
