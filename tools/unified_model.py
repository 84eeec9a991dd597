# Python model of the unified (probe/copy-tail) per-thread loop used by the thread-per-message kernel.
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
from compress_model import hashb, emit_literal, emit_copy, oracle_frag
M32=0xffffffff
def model(data):
    n=len(data); inp=bytes(data)+bytes(8)
    def ld32(p): return int.from_bytes(inp[p:p+4],'little')
    tsize=256
    while tsize<16384 and tsize<n: tsize<<=1
    shift=32-(tsize.bit_length()-1)
    table=[0]*tsize; out=[]; next_emit=0
    if n>=15:
        ip_limit=n-15; ip=1; probing=True; skip=32
        while True:
            if probing:
                step=skip>>5; skip+=1; nxt=ip+step
                if nxt>ip_limit: break
            cur=ld32(ip); h=hashb(cur,shift); cand=table[h]; table[h]=ip
            if ld32(cand)==cur:
                if ip>next_emit: emit_literal(out,inp[next_emit:ip],ip-next_emit)
                m=4
                while ip+m<n and inp[cand+m]==inp[ip+m]: m+=1
                emit_copy(out,ip-cand,m); ip+=m; next_emit=ip
                if ip>=ip_limit: break
                table[hashb(ld32(ip-1),shift)]=ip-1
                probing=False
            else:
                if probing: ip=nxt
                else: ip+=1; probing=True; skip=32
    if next_emit<n: emit_literal(out,inp[next_emit:n],n-next_emit)
    return bytes(out)
rng=np.random.default_rng(5); cases=[]
for kind in ('az09','rand62','words'):
    for n in (1,14,15,16,17,40,100,257,1000,1024,1025,4096,20000,65536):
        cases.append(O.corpus(kind,n,n).tobytes())
cases+= [bytes(65536), b'ab'*3000, bytes(300)+O.corpus('rand62',700,3).tobytes()+bytes(300)]
for t in range(150):
    n=int(rng.integers(1,3000)); cases.append(rng.integers(0,int(rng.integers(2,6)),n,dtype=np.uint8).tobytes())
for t in range(500):
    cases.append(O.corpus('words',int(rng.integers(1,1025)),t).tobytes())
bad=sum(model(d)!=oracle_frag(d) for d in cases)
print(len(cases),'cases bad',bad)
