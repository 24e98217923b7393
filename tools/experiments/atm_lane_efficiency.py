"""CPU model of the ATM kernel warp schedule: lane efficiency of the lane-per-edge mapping for different orderings of the
neighbour list (cell order, distance shells, Morton sub-cells), with and without CTA-wide lockstep per chunk.
Result on water3k (this round): 0.45-0.49 in lockstep, 0.53-0.63 free-running, whatever the ordering."""
import numpy as np, sys
sys.path.insert(0,'/root/repo')
from dftd4_b200 import workloads
w=workloads.get("water3k")
pos=w["positions"]; lat=w["lattice"]; n=len(pos)
rng=np.random.default_rng(0)
cut=40.0
T=np.array([[a,b,c] for a in (-1,0,1) for b in (-1,0,1) for c in (-1,0,1)])@lat
allp=(pos[None,:,:]+T[:,None,:]).reshape(-1,3)
atom=np.tile(np.arange(n),27)
def morton(ix,iy,iz,bits=4):
    m=np.zeros_like(ix)
    for b in range(bits):
        m|=((ix>>b)&1)<<(3*b)| ((iy>>b)&1)<<(3*b+1) | ((iz>>b)&1)<<(3*b+2)
    return m
COST_FULL=78; COST_SKIP=13
def simulate(i, order_kind, lockstep=True, KC=128, chunk_cull=False):
    d=allp-pos[i]; r2=(d*d).sum(1)
    sel=(r2<=cut*cut)&(r2>1e-10)
    P=d[sel]; A=atom[sel]; R=np.sqrt(r2[sel])
    # base order = cell order (20 bohr cells, z,y,x) then atom index
    gp=allp[sel]
    cell=np.floor((gp-allp.min(0))/20.0).astype(int)
    base=np.lexsort((A,cell[:,0],cell[:,1],cell[:,2]))
    if order_kind=="cell": o=base
    elif order_kind=="shell":
        sh=np.minimum(15,(R/cut*16).astype(int)); o=base[np.argsort(sh[base],kind="stable")]
    elif order_kind.startswith("morton"):
        h=float(order_kind[6:]); q=np.floor((P+cut)/h).astype(int); o=np.argsort(morton(q[:,0],q[:,1],q[:,2]),kind="stable")
    elif order_kind.startswith("shmorton"):
        h=float(order_kind[8:]); q=np.floor((P+cut)/h).astype(int); m=morton(q[:,0],q[:,1],q[:,2]); sh=np.minimum(3,(R/cut*4).astype(int)); o=np.lexsort((m,sh))
    P=P[o]; A=A[o]
    own=np.where(A<=i)[0]
    nk=len(P)
    tot=0.0; ideal=0.0; acc=0
    for p0 in range(0,len(own),256):
        ed=own[p0:p0+256]
        J=P[ed]
        D=((J[:,None,:]-P[None,:,:])**2).sum(2)   # edges x k
        ok=(D<=cut*cut)&(D>1e-10)
        acc+=ok.sum()
        nw=(len(ed)+31)//32
        anyw=np.zeros((nw,nk),bool)
        for wi in range(nw): anyw[wi]=ok[wi*32:(wi+1)*32].any(0)
        for c0 in range(0,nk,KC):
            a=anyw[:,c0:c0+KC].sum(1); kc=min(KC,nk-c0)
            cost=a*COST_FULL+(kc-a)*COST_SKIP
            if chunk_cull: cost=np.where(a==0,0,cost)
            # 8 warps over 4 SMSPs: lockstep => time = max per chunk; count in warp-instr units *8 warps
            tot+= (cost.max()*8 if lockstep else cost.sum()+ (8-nw)*0)
    ideal=acc/32*COST_FULL
    return ideal,tot,acc
cent=rng.choice(n,12,replace=False)
cent=np.concatenate([cent,[n-1,n-5, n//2]])
for kind in ["cell","shell","morton10","morton5","shmorton10","morton7"]:
    for lock in (True,False):
        for cc in (False,True):
            I=Tt=0
            for i in cent:
                a,b,_=simulate(i,kind,lock,128,cc); I+=a;Tt+=b
            print(kind,"lockstep" if lock else "free","cull" if cc else "", "efficiency %.3f"%(I/Tt))
