import sys, os, hashlib
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/oracle'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
import bn256_ref as R
from drynx_b200 import Backend
from conftest import SplitMix, Oracle
be=Backend(0); rng=SplitMix(5); orc=Oracle()
a=np.frombuffer(bytes(range(32)),np.uint32).copy(); b=np.frombuffer(bytes(range(32,64)),np.uint32).copy()
d=be.mont_mul(6,a,b).tobytes()+be.mont_mul(7,a,b).tobytes()
print('sha3 dev ok', d==hashlib.sha3_512(bytes(range(64))).digest())
d=be.mont_mul(8,a,b).tobytes()+be.mont_mul(9,a,b).tobytes()
print('sha3 3-block dev ok', d==hashlib.sha3_512(bytes(range(64))*3).digest())
sk=rng.scalar(); Pp=R.g1_mul(sk,R.G1_GEN); P=R.g1_to_bytes(Pp); be.set_pubkey(P)
ns,u,l=1,2,1
xs=[rng.scalar()]
Y,A=orc.make_sigs(xs,u)
be.set_signatures(0,ns,u,Y,A)
r=rng.scalar(); ct=R.encrypt(Pp,1,r)
rnd=rng.scalars_bytes(l*(3+ns))
pr,off=be.create_range_proofs([0],{0:ns},{0:u},[l],[1],r.to_bytes(32,'big'),ct,[rnd])
want,_=orc.rp_create(ns,u,l,Y,A,P,[1],r.to_bytes(32,'big'),ct,[rnd])
pr=pr.tobytes()
names=[('ct',0,128),('c',128,160),('Zr',160,192),('D',192,256),('Zphi',256,288),('V',288,416),('A',416,800),('Zv',800,832)]
for nm,a0,a1 in names: print(nm, pr[a0:a1]==want[a0:a1])
c_def=int.from_bytes(R.sha3_512(R.g1_to_bytes(R.G1_GEN)+ct[64:128]+Y),'big')%R.n
print('c dev', pr[128:160].hex()); print('c def', c_def.to_bytes(32,'big').hex()); print('c orc', want[128:160].hex())

d=R.sha3_512(R.g1_to_bytes(R.G1_GEN)+ct[64:128]+Y)
import struct
m=R.g1_to_bytes(R.G1_GEN)+ct[64:128]+Y
w=struct.unpack('<48I',m); print('want msg %08x %08x | %08x %08x | %08x %08x'%(w[0],w[15],w[16],w[31],w[32],w[47]))
dw=struct.unpack('<16I',d); print('want dig %08x %08x %08x %08x'%(dw[0],dw[7],dw[8],dw[15]))
print('want c   %08x %08x'%(c_def&0xffffffff, c_def>>224))
