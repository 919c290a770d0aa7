"""CPU tool: the bit manipulation of the N4 weight-unit loaders of csrc/mmvq.cu (Q2_K, Q3_K, IQ4_NL, IQ4_XS, MXFP4, NVFP4), restated with Python integers
(PRMT and the saturating byte subtract emulated), against dequantize_row_* of the unmodified reference (oracle/_ref) — checks the format
reading before GPU time is spent.  Test infrastructure only."""
import numpy as np, struct, sys
sys.path.insert(0,'/root/repo')
from oracle import refenv
ref = refenv.load_ref()          # None when oracle/_ref was not built
M32=0xFFFFFFFF
def f16(b0,b1): return float(np.frombuffer(bytes([b0,b1]),dtype=np.float16)[0])
def word(b,off): return int.from_bytes(bytes(b[off:off+4]),'little')
def prmt(a,b,sel):
    bs = list(a.to_bytes(4,'little'))+list(b.to_bytes(4,'little'))
    out=0
    for i in range(4):
        n=(sel>>(4*i))&0xF
        assert n<8
        out |= bs[n]<<(8*i)
    return out
def pack_s8(a,b,c,d): return (a&0xFF)|((b&0xFF)<<8)|((c&0xFF)<<16)|((d&0xFF)<<24)
IQ=[pack_s8(-127,-104,-83,-65),pack_s8(-49,-35,-22,-10),pack_s8(1,13,25,38),pack_s8(53,69,89,113)]
MX=[pack_s8(0,1,2,3),pack_s8(4,6,8,12),pack_s8(0,-1,-2,-3),pack_s8(-4,-6,-8,-12)]
def tab16(x,T):
    y=x&0x07070707
    sel=prmt((y|(y>>4))&M32,0,0x4420)
    lo=prmt(T[0],T[1],sel&0xFFFF); hi=prmt(T[2],T[3],sel&0xFFFF)
    top=(((x>>3)&0x01010101)*0xFF)&M32
    return (hi&top)|(lo&~top&M32)
def s8x4(w): return [ (b-256 if b>127 else b) for b in w.to_bytes(4,'little')]
def vsubss4(a,b):
    return [max(-128,min(127,x-y)) for x,y in zip(s8x4(a),s8x4(b))]
def e8m0_half(e):
    bits = (0x00200000<<e) if e<2 else ((e-1)<<23)
    return struct.unpack('f',struct.pack('I',bits))[0]
def ue4m3_half(x):
    if x==0 or x==0x7F: return 0.0
    e=(x>>3)&0xF; man=x&7
    if e==0: return man*2.0**-10
    return struct.unpack('f',struct.pack('I',((e+119)<<23)|(man<<20)))[0]

def unit_q2k(row,ib):
    sb,j=ib>>3,ib&7; n,r=j>>2,j&3
    b=row[sb*84:(sb+1)*84]
    sc=(word(b,4*(2*n+(r>>1)))>>(16*(r&1)))&0xFFFF
    d,dm=f16(b[80],b[81]),f16(b[82],b[83])
    d0,m0,d1,m1=d*(sc&0xF),dm*((sc>>4)&0xF),d*((sc>>8)&0xF),dm*(sc>>12)
    out=[]
    for i in range(8):
        v=(word(b,4*(4+8*n+i))>>(2*r))&0x03030303
        for q in v.to_bytes(4,'little'): out.append((d0 if i<4 else d1)*q-(m0 if i<4 else m1))
    return out
def unit_q3k(row,ib):
    sb,j=ib>>3,ib&7; n,r=j>>2,j&3
    b=row[sb*110:(sb+1)*110]
    raw=b[96:108]; is_=2*j; sh=2*(is_>>2)
    b0,b1=raw[is_&7],raw[(is_&7)+1]
    l0=(b0&0xF) if j<4 else (b0>>4); l1=(b1&0xF) if j<4 else (b1>>4)
    h0=(raw[8+(is_&3)]>>sh)&3; h1=(raw[9+(is_&3)]>>sh)&3
    d=f16(b[108],b[109]); d0=d*((l0|(h0<<4))-32); d1=d*((l1|(h1<<4))-32)
    out=[]
    for i in range(8):
        lw=word(b,32+32*n+4*i); hw=word(b,4*i)
        v=((lw>>(2*r))&0x03030303)|(((hw>>j)&0x01010101)<<2)
        for q in vsubss4(v,0x04040404): out.append((d0 if i<4 else d1)*q)
    return out
def unit_nib(words,T,d):
    lo=[];hi=[]
    for w in words:
        lo+=s8x4(tab16(w&0x0F0F0F0F,T)); hi+=s8x4(tab16((w>>4)&0x0F0F0F0F,T))
    return [d*x for x in lo+hi]
def unit_iq4nl(row,ib):
    b=row[ib*18:(ib+1)*18]
    return unit_nib([word(b,2+4*i) for i in range(4)],IQ,f16(b[0],b[1]))
def unit_iq4xs(row,ib):
    sb,j=ib>>3,ib&7
    b=row[sb*136:(sb+1)*136]
    w0,sl=word(b,0),word(b,4)
    ls=((sl>>(4*j))&0xF)|((((w0>>16)>>(2*j))&3)<<4)
    d=f16(b[0],b[1])*(ls-32)
    return unit_nib([word(b,4*(2+4*j+i)) for i in range(4)],IQ,d)
def unit_mxfp4(row,ib):
    b=row[ib*17:(ib+1)*17]
    return unit_nib([word(b,1+4*i) for i in range(4)],MX,e8m0_half(b[0]))
def unit_nvfp4(row,ib):
    b=row[(ib>>1)*36:((ib>>1)+1)*36]; h=ib&1
    sc=word(b,0)>>(16*h)
    d0,d1=ue4m3_half(sc&0xFF),ue4m3_half((sc>>8)&0xFF)
    out=[]
    for s in range(2):
        w0,w1=word(b,4*(1+4*h+2*s)),word(b,4*(2+4*h+2*s))
        v=[tab16(w0&0x0F0F0F0F,MX),tab16(w1&0x0F0F0F0F,MX),tab16((w0>>4)&0x0F0F0F0F,MX),tab16((w1>>4)&0x0F0F0F0F,MX)]
        for x in v: out+=[(d0 if s==0 else d1)*q for q in s8x4(x)]
    return out
TYPES={"q2_K":(10,256,84,unit_q2k),"q3_K":(11,256,110,unit_q3k),"iq4_nl":(20,32,18,unit_iq4nl),"iq4_xs":(23,256,136,unit_iq4xs),"mxfp4":(39,32,17,unit_mxfp4),"nvfp4":(40,64,36,unit_nvfp4)}
def check(name, k=1024, seed=1):
    """max |emulated - dequantize_row| / max |dequantize_row| for one row of gaussian weights quantized by the reference"""
    t,bs,sz,fn=TYPES[name]
    rng=np.random.default_rng(seed)
    wf=(rng.standard_normal((2,k))*0.05).astype(np.float32)
    W=np.zeros((2,k//bs*sz),np.uint8)
    ref.ref_quantize_chunk(t,refenv.ptr(wf),refenv.ptr(W),2,k)
    want=np.zeros(k,np.float32)
    assert ref.ref_dequantize_row(t,refenv.ptr(np.ascontiguousarray(W[1])),refenv.ptr(want),k)==0
    row=[int(x) for x in W[1]]
    got=np.array(sum([fn(row,ib) for ib in range(k//32)],[]),np.float64)
    return float(np.abs(got-want).max()/np.abs(want).max()), int(np.count_nonzero(want))

if __name__=="__main__":
    for name in TYPES:
        err,nz=check(name)
        print(name,"max rel err",err,"nonzero",nz)
