"""Round 2, experiment d (CPU, numpy model of the kernel, tests/ipm_model.py): one step of iterative refinement of the corrector
solve -- residual r = rhs - (Dblk + W^T M W) du evaluated in FP64, correction solved with the same stage factors -- on random
problems at N = 40 and N = 64.  Result (profiles/r2_d_refinement_model.txt): it does NOT lower the dual-residual floor and costs
convergence (several problems run into max_iter).  The stage-wise solve is already backward stable to the level at which the
residual itself can be evaluated in FP64 (|K| |du| eps with |K| ~ N^3 dt^2 Q); fixed-precision refinement has nothing to gain there.
The remedy VERDICT r1 asked for was therefore not built; long horizons report G1MPC_DUAL_RESIDUAL where the floor exceeds dual_tol."""
import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import ipm_model as im
from g1_locomotion_b200 import presets, inputs
import torch

def solve(prob, refine, linear="stage", gap_tol=1e-10, max_iter=40):
    N, nc, mu, sr = prob.N, prob.nc, prob.mu, prob.sr
    M, h = prob.M_h()
    st = prob.stance
    Ct = lambda q: np.array([q[0]-q[1], q[2]-q[3], -mu*(q[0]+q[1]+q[2]+q[3]) + q[4]-q[5]])
    Cb = np.array([[1,0,-mu],[-1,0,-mu],[0,1,-mu],[0,-1,-mu],[0,0,1],[0,0,-1.0]])
    u = np.zeros((N,nc,3)); s = np.ones((N,nc,6)); lam = np.zeros((N,nc,6))
    for k in range(N):
        ns = int(st[k].sum())
        for j in range(nc):
            if st[k,j]:
                span = prob.hi[k,j]-prob.lo[k,j]
                fz = min(max(prob.weight/ns, prob.lo[k,j]+0.1*span), prob.hi[k,j]-0.1*span)
                u[k,j,2]=fz; s[k,j]=[mu*fz]*4+[prob.hi[k,j]-fz, fz-prob.lo[k,j]]
    m_ineq = 6.0*max(int(st.sum()),1)
    Wk=[prob.W(k) for k in range(N)]
    def grad(u, with_h=True):
        w=np.array([Wk[k]@u[k].ravel() for k in range(N)])
        v=(M@w.ravel()+(h if with_h else 0)).reshape(N,6)
        return np.array([sr*u[k].ravel()+Wk[k].T@v[k] for k in range(N)]).reshape(N,nc,3)
    gf=grad(u); g0=np.abs(grad(np.zeros_like(u))).max()
    mu0=max(1e-2,0.1*np.abs(gf[st]).max()); lam[st]=mu0/s[st]
    it=0
    while True:
        gf=grad(u)
        gap=float((s[st]*lam[st]).sum()/m_ineq)
        rd=max(np.abs(gf[k,j]+Ct(lam[k,j])).max() for k in range(N) for j in range(nc) if st[k,j])
        if gap<=gap_tol or it>=max_iter:
            return u,it,gap,rd/(1+g0)
        Dinv=np.zeros((N,nc,3,3)); X=np.zeros((N,nc,6,3)); S=np.zeros((N,6,6)); Dm=np.zeros((N,nc,3,3))
        for k in range(N):
            for j in range(nc):
                if st[k,j]:
                    D=lam[k,j]/s[k,j]
                    Dinv[k,j],X[k,j]=im._contact_blocks(D,sr,mu)
                    Dm[k,j]=sr*np.eye(3)+Cb.T@(D[:,None]*Cb)
                    Wc=Wk[k][:,3*j:3*j+3]; S[k]+=Wc@Dinv[k,j]@Wc.T
        fac=im.StageFactor(prob,S,theta=True)
        def newton(rhs):
            tw=np.zeros((N,6))
            for k in range(N):
                for j in range(nc):
                    if st[k,j]: tw[k]+=Wk[k][:,3*j:3*j+3]@(Dinv[k,j]@rhs[k,j])
            y=fac.solve(tw)
            du=np.zeros_like(u); wy=np.zeros_like(u)
            for k in range(N):
                for j in range(nc):
                    if st[k,j]:
                        wy[k,j]=Wk[k][:,3*j:3*j+3].T@y[k]
                        du[k,j]=Dinv[k,j]@(rhs[k,j]-wy[k,j])
            return du,wy
        ds=np.zeros_like(s); dl=np.zeros_like(s)
        sigma_gap=0.0
        for pas in range(2):
            q=np.zeros_like(s)
            if pas==1: q[st]=(min(1.0,a_aff/0.3)*ds[st]*dl[st]-sigma_gap)/s[st]
            rhs=np.zeros_like(u)
            for k in range(N):
                for j in range(nc):
                    if st[k,j]: rhs[k,j]=-gf[k,j]+Ct(q[k,j])
            du,wy=newton(rhs)
            if refine and pas==1:
                for _ in range(refine):
                    Kdu=grad(du,with_h=False)-sr*du   # W'MW du
                    r=np.zeros_like(u)
                    for k in range(N):
                        for j in range(nc):
                            if st[k,j]: r[k,j]=rhs[k,j]-Dm[k,j]@du[k,j]-Kdu[k,j]
                    d2,wy2=newton(r)
                    du=du+d2; wy=wy+wy2
            amin=1e300
            for k in range(N):
                for j in range(nc):
                    if st[k,j]:
                        rt=rhs[k,j]-wy[k,j]
                        ds[k,j]=-(X[k,j]@rt)
                        dl[k,j]=-lam[k,j]-q[k,j]-(lam[k,j]/s[k,j])*ds[k,j]
                        for i in range(6):
                            if ds[k,j,i]<0: amin=min(amin,-s[k,j,i]/ds[k,j,i])
                            if dl[k,j,i]<0: amin=min(amin,-lam[k,j,i]/dl[k,j,i])
            if pas==0:
                a_aff=min(1.0,amin); g_aff=float(((s[st]+a_aff*ds[st])*(lam[st]+a_aff*dl[st])).sum()/m_ineq)
                sig=(g_aff/gap)**3; sigma_gap=sig*gap
            else:
                alpha=min(1.0,min(0.99999,max(0.995,1.0-(1.0-a_aff)**2))*amin)
        u[st]+=alpha*du[st]; lam[st]+=alpha*dl[st]; s[st]+=alpha*ds[st]
        it+=1

for N in (40, 64):
    pr=presets.get_preset("cpp_node",horizon=N)
    b=[t.numpy() for t in inputs.random_batch(pr,6,seed=17)]
    for i in range(6):
        P=im.Problem(pr,*[a[i] for a in b])
        r0=solve(P,0); r1=solve(P,1)
        print(N,i,"no refine: it %d rd %.1e | refine: it %d rd %.1e | du %.1e"%(r0[1],r0[3],r1[1],r1[3],np.abs(r0[0]-r1[0]).max()))
