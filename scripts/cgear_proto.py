"""Go / no-go prototype for round 2: single-reduction (Chronopoulos-Gear) PCG against the standard two-reduction PCG the
library runs, both with the trilinear V(1,1) preconditioner, vectors stored in fp32 and dot products in fp64 (as cg.cuh does).
One all-reduce per iteration instead of two is what the multi-GPU iteration needs.

    python scripts/cgear_proto.py
    (32, 64, 32)   rtol 1e-06: standard 10 iterations (true residual 2.8e-07), single-reduction 10 (2.8e-07)
    (64, 128, 64)  rtol 1e-07: standard 12 (1.3e-07), single-reduction 12 (1.3e-07)
Same counts, same residuals: the recurrence for <d, M d> costs nothing in fp32 at these iteration counts.
Cost on the GPU: the update pass reads u, w, d, s, x, r and writes d, s, x, r (48 B/cell with the matvec, against 40 today).
"""
import sys, numpy as np
sys.argv = ['x']; sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.abspath(__file__)))
import mg_proto as P
def M32(x): return P.M(x.astype(np.float64)).astype(np.float32)
def pcg_std(f, B, rtol=1e-6, maxit=60, dt=np.float32):
    f=f.astype(dt); x=np.zeros_like(f); r=f.copy(); z=B(r).astype(dt); d=z.copy()
    rz=float(np.vdot(r.astype(np.float64),z)); bb=float(np.vdot(f.astype(np.float64),f))
    for it in range(1,maxit+1):
        w=M32(d) if dt==np.float32 else P.M(d)
        a=rz/float(np.vdot(d.astype(np.float64),w))
        x=(x+dt(a)*d).astype(dt); r=(r-dt(a)*w).astype(dt)
        rr=float(np.vdot(r.astype(np.float64),r))
        if rr<=rtol*rtol*bb: return x,it
        z=B(r).astype(dt); rzn=float(np.vdot(r.astype(np.float64),z)); d=(z+dt(rzn/rz)*d).astype(dt); rz=rzn
    return x,maxit
def pcg_cg1(f, B, rtol=1e-6, maxit=60, dt=np.float32):
    # Chronopoulos-Gear: one reduction (gamma=<r,u>, delta=<w,u>, rr) per iteration
    f=f.astype(dt); x=np.zeros_like(f); r=f.copy(); bb=float(np.vdot(f.astype(np.float64),f))
    u=B(r).astype(dt); w=M32(u) if dt==np.float32 else P.M(u)
    gamma=float(np.vdot(r.astype(np.float64),u)); delta=float(np.vdot(w.astype(np.float64),u))
    alpha=gamma/delta; beta=0.0
    d=np.zeros_like(f); s=np.zeros_like(f)
    for it in range(1,maxit+1):
        d=(u+dt(beta)*d).astype(dt); s=(w+dt(beta)*s).astype(dt)
        x=(x+dt(alpha)*d).astype(dt); r=(r-dt(alpha)*s).astype(dt)
        u=B(r).astype(dt); w=M32(u) if dt==np.float32 else P.M(u)
        gn=float(np.vdot(r.astype(np.float64),u)); delta=float(np.vdot(w.astype(np.float64),u)); rr=float(np.vdot(r.astype(np.float64),r))
        if rr<=rtol*rtol*bb: return x,it
        beta=gn/gamma; alpha=gn/(delta-beta*gn/alpha); gamma=gn
    return x,maxit
for shape in [(32,64,32),(64,128,64)]:
    rng=np.random.default_rng(0); u=rng.standard_normal((3,)+shape)
    f=np.diff(u[0],axis=0,append=0)+np.diff(u[1],axis=1,append=0)+np.diff(u[2],axis=2,append=0)
    mg=P.MG(shape,'tri',omega=0.8,nu=1,coarse_cells=2048)
    B=lambda r: mg.cycle(r.astype(np.float64))
    for rtol in (1e-6,1e-7):
        xs,i1=pcg_std(f,B,rtol); xc,i2=pcg_cg1(f,B,rtol)
        res=lambda x: np.linalg.norm(f-P.M(x.astype(np.float64)))/np.linalg.norm(f)
        print(shape,rtol,'std',i1,'%.2e'%res(xs),'cgear',i2,'%.2e'%res(xc))
