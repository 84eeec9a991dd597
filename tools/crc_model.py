# numeric model of the quad-per-chunk CRC32C decomposition (validates the algebra before CUDA)
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
POLY=0x82F63B78
T0=[0]*256
for i in range(256):
    c=i
    for _ in range(8): c=(c>>1)^(POLY if c&1 else 0)
    T0[i]=c
def Z1(l): return T0[l&0xff]^(l>>8)
def Z(l,k):
    for _ in range(k): l=Z1(l)
    return l
def mm(a,b):  # reflected multiply mod P: bit31 = x^0
    p=0
    for i in range(32):
        if (a>>(31-i))&1: p^=b
        b=(b>>1)^(POLY if b&1 else 0)
    return p
ONE=1<<31
def xpow8(k):  # x^(8k) mod P
    r=ONE
    X8=Z(ONE,1)  # ? check: Z1 of ONE = x^8 ?
    # square-multiply
    base=X8; 
    while k:
        if k&1: r=mm(r,base)
        base=mm(base,base); k>>=1
    return r
import random
random.seed(1)
for t in range(20):
    l=random.getrandbits(32); k=random.randrange(0,200)
    assert Z(l,k)==mm(l,xpow8(k)),(t,)
print("Z^k(l)==mm(l,x^(8k)) ok; X8=",hex(Z(ONE,1)))
# S(l,data)
def S(l,data):
    for b in data: l=T0[(l^b)&0xff]^(l>>8)
    return l
p=O.port()
def crc_quad(data,init=0):
    n=len(data); l0=init^0xffffffff
    if n<4:
        return S(l0,data)^0xffffffff
    rows=n//64
    w=np.frombuffer(bytes(data)+bytes(64+(-len(data))%4),dtype='<u4')
    H=[[0]*4 for _ in range(4)]
    for it in range(rows):
        for q in range(4):
            for c in range(4):
                x=int(w[it*16+q*4+c])
                if it==0 and q==0 and c==0: x^=l0
                H[q][c]=Z(H[q][c],64)^x if it>0 else x
    if rows>0:
        G=[0]*4
        for q in range(4):
            g=0
            for c in range(4): g=Z(g^H[q][c],4)
            G[q]=g
        P01=Z(G[0],16)^G[1]; P23=Z(G[2],16)^G[3]
        Rm=Z(P01,32)^P23
    else: Rm=0
    t=n-rows*64; tw=t//4; tb=t%4
    acc=0
    for j in range(16):
        if j<tw:
            x=int(w[rows*16+j])
            if j==0:
                x^=Rm
                if rows==0: x^=l0
            acc^=mm(x,xpow8(4*(tw-j)))
    if tw==0: acc=Rm   # (rows>0 guaranteed since n>=4... no: n>=4 and rows==0 -> tw>=1)
    tail=data[rows*64+tw*4:]
    acc=S(acc,tail)
    return acc^0xffffffff
rng=np.random.default_rng(0)
for n in list(range(0,200))+[1023,1024,1025,4096,4097,5000]:
    d=rng.integers(0,256,n,dtype=np.uint8)
    init=int(rng.integers(0,2**32))
    assert crc_quad(d.tobytes(),init)==p.crc32c_extend(init,d),n
print("quad model ok")
# multi-chunk combine
CH=4096
def crc_chunks(data,init=0):
    n=len(data); l0=init^0xffffffff
    K=max(1,(n+CH-1)//CH)
    rem=[]
    for k in range(K):
        ch=data[k*CH:(k+1)*CH]
        # raw S(l0 or 0, chunk): use crc_quad with init trick: crc_quad returns S(init^~0)^~0
        r=crc_quad(ch,(0xffffffff if k>0 else init))^0xffffffff   # init=~0 -> l0=0
        rem.append(r)
    acc=rem[0]
    for k in range(1,K):
        lk=len(data[k*CH:(k+1)*CH])
        acc=mm(acc,xpow8(lk))^rem[k]
    return acc^0xffffffff
for n in [4096,4097,8191,8192,8193,20000,70001]:
    d=rng.integers(0,256,n,dtype=np.uint8); init=int(rng.integers(0,2**32))
    assert crc_chunks(d.tobytes(),init)==p.crc32c_extend(init,d),n
print("chunk combine ok")
