# Join ncu per-SASS-instruction counts with nvdisasm line info -> per-source-line instruction counts.
import csv, re, subprocess, sys, collections
cubin, kernel_pat, ncu_csv, srcfile = sys.argv[1:5]
dis = subprocess.run(['nvdisasm','-g','-c',cubin],capture_output=True,text=True).stdout.splitlines()
# find function section
line_of = {}  # offset -> (file,line)
cur=None; infn=False
for l in dis:
    if l.startswith('.text.'):
        infn = kernel_pat in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur=(m.group(1).split('/')[-1], int(m.group(2)))
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m and infn:
        line_of[int(m.group(1),16)] = (cur, m.group(2))
rows=list(csv.reader(open(ncu_csv)))
hdr=rows[1]; ix={h:i for i,h in enumerate(hdr)}
base=None; agg=collections.Counter(); tot=0; samples=collections.Counter()
for r in rows[2:]:
    if len(r)<len(hdr): continue
    a=int(r[ix['Address']],16)
    if base is None: base=a
    off=a-base
    n=int(r[ix['Instructions Executed']]); s=int(r[ix['# Samples']])
    key=line_of.get(off,((None,0),''))[0]
    agg[key]+=n; samples[key]+=s; tot+=n
src={}
try:
    for i,l in enumerate(open(srcfile),1): src[i]=l.rstrip()
except Exception: pass
print('total inst',tot)
for key,n in agg.most_common(45):
    f,ln=key if key else (None,0)
    print(f"{n/tot*100:5.1f}% smp={samples[key]:6d} {f}:{ln}: {src.get(ln,'')[:110] if f and f in srcfile else ''}")
