# Lane-level Python model of compress_fragment_warp (mirrors snappy_compress.cu) checked against the oracle.
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
M32=0xffffffff
def hashb(b,shift): return ((b*0x1e35a7bd)&M32)>>shift
def emit_literal(out,src,ln):
    n=ln-1
    if n<60: out.append(n<<2)
    else:
        cnt=1 if n<256 else 2 if n<65536 else 3 if n<(1<<24) else 4
        out.append((59+cnt)<<2)
        for i in range(cnt): out.append((n>>(8*i))&0xff)
    out.extend(src)
def emit_copy(out,offset,ln):
    n64=(ln-68)//64+1 if ln>=68 else 0
    for k in range(n64): out.extend([2+(63<<2),offset&0xff,offset>>8])
    ln-=64*n64
    if ln>64:
        out.extend([2+(59<<2),offset&0xff,offset>>8]); ln-=60
    if ln<12 and offset<2048: out.extend([1+((ln-4)<<2)+((offset>>8)<<5),offset&0xff])
    else: out.extend([2+((ln-1)<<2),offset&0xff,offset>>8])
def model(data):
    n=len(data); inp=bytes(data)+bytes(8)
    def ld32(p): return int.from_bytes(inp[p:p+4],'little')
    tsize=256
    while tsize<16384 and tsize<n: tsize<<=1
    shift=32-(tsize.bit_length()-1)
    table=[0]*tsize
    out=[]; next_emit=0
    if n>=15:
        ip_limit=n-15; ip=1
        while True:
            base=ip; s=1; found=False; cand=0
            while True:
                pos=[base+l*s for l in range(32)]
                valid=[p+s<=ip_limit for p in pos]
                cur=[ld32(pos[l]) if valid[l] else 0 for l in range(32)]
                h=[hashb(cur[l],shift) if valid[l] else (0x80000000|l) for l in range(32)]
                peers=[sum(1<<j for j in range(32) if h[j]==h[l]) for l in range(32)]
                c=[0]*32; match=[False]*32
                for l in range(32):
                    if valid[l]:
                        prev=peers[l]&((1<<l)-1)
                        c[l]=base+(prev.bit_length()-1)*s if prev else table[h[l]]
                        match[l]= ld32(c[l])==cur[l]
                mball=sum(1<<l for l in range(32) if match[l]); vball=sum(1<<l for l in range(32) if valid[l])
                fi=32 if vball==M32 else ((~vball&M32)&-(~vball&M32)).bit_length()-1
                fm=(mball&-mball).bit_length()-1 if mball else 32
                reach=fm+1 if fm<fi else fi
                commit=M32 if reach>=32 else (1<<reach)-1
                for l in range(32):
                    if (commit>>l)&1 and (peers[l]&commit&~((1<<l)-1)&~(1<<l)&M32)==0:
                        table[h[l]]=pos[l]&0xffff
                if fm<fi:
                    found=True; ip=base+fm*s; cand=c[fm]; break
                if fi<32: break
                base+=32*s; s+=1
            if not found: break
            emit_literal(out,inp[next_emit:ip],ip-next_emit)
            at_end=False
            while True:
                matched=0
                while True:
                    nbs=[]
                    for l in range(32):
                        p2=ip+matched+4*l; p1=cand+matched+4*l
                        vb=min(4,n-p2) if p2<n else 0
                        nb=0
                        if vb:
                            x=ld32(p1)^ld32(p2)
                            nb=((x&-x).bit_length()-1)>>3 if x else 4
                            nb=min(nb,vb)
                        nbs.append(nb)
                    stop=[l for l in range(32) if nbs[l]<4]
                    if stop:
                        f=stop[0]; matched+=4*f+nbs[f]; break
                    matched+=128
                if matched<4: break
                emit_copy(out,ip-cand,matched)
                ip+=matched; next_emit=ip
                if ip>=ip_limit: at_end=True; break
                w0=ld32(ip-1); w1=ld32(ip)
                hp=hashb(w0,shift); hc=hashb(w1,shift)
                table[hp]=(ip-1)&0xffff; cand=table[hc]; table[hc]=ip&0xffff
            if at_end: break
            ip+=1
    if next_emit<n: emit_literal(out,inp[next_emit:n],n-next_emit)
    return bytes(out)
p=O.port(); lib=p.lib
def oracle_frag(d):
    a=np.frombuffer(bytes(d),dtype=np.uint8) if len(d) else np.zeros(1,dtype=np.uint8)
    out=np.empty(32+len(d)+len(d)//6+16,dtype=np.uint8)
    k=lib.b2o_snappy_compress_fragment(a.ctypes.data_as(O._u8p),len(d),out.ctypes.data_as(O._u8p))
    return out[:k].tobytes()
rng=np.random.default_rng(5)
cases=[]
for kind in ('az09','rand62','words'):
    for n in (1,14,15,16,17,40,100,257,1000,1024,1025,4096,5000,20000,65536):
        cases.append(O.corpus(kind,n,n).tobytes())
cases.append(bytes(65536)); cases.append(b'ab'*3000); cases.append(bytes(300)+O.corpus('rand62',700,3).tobytes()+bytes(300))
for t in range(60):
    n=int(rng.integers(1,6000)); cases.append(rng.integers(0,3,n,dtype=np.uint8).tobytes())
for t in range(20):
    n=int(rng.integers(1,3000)); cases.append(rng.integers(0,256,n,dtype=np.uint8).tobytes())
bad=0
for i,d in enumerate(cases):
    a=model(d); b=oracle_frag(d)
    if a!=b:
        bad+=1; print("MISMATCH case",i,len(d),len(a),len(b))
print("cases",len(cases),"bad",bad)
