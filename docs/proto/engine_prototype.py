"""numpy prototype of the kernel's algorithm in slack-eliminated stage-output form."""
import numpy as np

class QP:
    """condensed data for one agent QP"""
    def __init__(s, P, a, N, cnt):
        # P: params dict-like (cfg), a: oracle assemble_dense dict
        s.N=N; s.cnt=cnt; nu=2*N
        Ak,Bk = a['Ak'],a['Bk']
        Q=np.array(P['Q']); s.Qs=np.array(P['Qs']); R=np.array(P['R']); dR=np.array(P['dR'])
        m_in=(8+cnt)*N
        x0=a['u'][m_in:m_in+9]; uold=a['u'][m_in+12*(N+1):m_in+12*(N+1)+2]
        q=a['q']; p=q[:12*(N+1)].reshape(N+1,12)[:,:9]/2
        Gam=np.zeros((N+1,9,nu)); c=np.zeros((N+1,9)); c[0]=x0
        for t in range(N):
            Gam[t+1]=Ak[t]@Gam[t]; Gam[t+1][:,2*t:2*t+2]+=Bk[t]; c[t+1]=Ak[t]@c[t]
        H=np.zeros((nu,nu)); g=np.zeros(nu)
        for t in range(1,N+1):
            H+=2*Gam[t].T@Q@Gam[t]; g+=2*Gam[t].T@(Q@c[t]+p[t])
        for i in range(N):
            # du_i = u_i - u_{i-1}
            H[2*i:2*i+2,2*i:2*i+2]+=2*(R+dR)
            if i>0:
                H[2*i-2:2*i,2*i-2:2*i]+=2*dR; H[2*i:2*i+2,2*i-2:2*i]-=2*dR; H[2*i-2:2*i,2*i:2*i+2]-=2*dR
        g[:2]+=-2*dR@uold
        s.H=H; s.g=g; s.Gam=Gam; s.c=c; s.x0=x0; s.uold=uold; s.p=p; s.Q=Q; s.R=R; s.dR=dR
        rows=[0,3,7,8]
        s.Gb=Gam[1:,rows,:]      # (N,4,nu)
        s.cb=c[1:,rows]          # (N,4)
        b=a['u'][:m_in]
        s.hs=b[:(4+cnt)*N].reshape(N,4+cnt)   # per stage rhs: -vmin, vmax, eb, eb, h_i...
        s.hb=b[(4+cnt)*N:].reshape(N,4)       # box rhs: max_rs,max_ls,max_ac,max_dc
        pl=a['planes']                        # (N,3,nn)
        s.ax=pl[:,0,:cnt]; s.ay=pl[:,1,:cnt]
    # constraint values r = A v - h (<=0 feasible) for v=(u,sig)
    def outputs(s,u): return np.einsum('tij,j->ti',s.Gb,u)+s.cb
    def resid(s,u,sig):
        o=s.outputs(u); N=s.N
        rs=np.zeros((N,4+s.cnt))
        rs[:,0]=-o[:,0]; rs[:,1]=o[:,0]+sig[:,0]; rs[:,2]=o[:,1]+sig[:,1]; rs[:,3]=-o[:,1]+sig[:,1]
        rs[:,4:]=s.ax*o[:,2:3]+s.ay*o[:,3:4]-sig[:,2:3]
        rs-=s.hs
        U=u.reshape(N,2)
        rb=np.stack([U[:,0],-U[:,0],U[:,1],-U[:,1]],1)-s.hb
        return rs,rb
    def grad_terms(s,ls,lb):
        """A^T lambda split: u-part (nu), sigma part (N,3)"""
        N=s.N
        wo=np.zeros((N,4))
        wo[:,0]=-ls[:,0]+ls[:,1]; wo[:,1]=ls[:,2]-ls[:,3]
        wo[:,2]=(s.ax*ls[:,4:]).sum(1); wo[:,3]=(s.ay*ls[:,4:]).sum(1)
        gu=np.einsum('tij,ti->j',s.Gb,wo)
        gu+= np.stack([lb[:,0]-lb[:,1], lb[:,2]-lb[:,3]],1).ravel()
        gs=np.stack([ls[:,1], ls[:,2]+ls[:,3], -ls[:,4:].sum(1)],1)
        return gu,gs
    def solve_weighted(s, Ds, Db, ru, rsig, extra_reg=0.0):
        """solve [H_u + A_u'DA_u, A_u'DA_s; A_s'DA_u, 2Qs + A_s'DA_s][du;dsig]=[ru;rsig] with row weights Ds (N,4+cnt), Db (N,4)"""
        N=s.N; cnt=s.cnt; nu=2*N
        # per stage 4x4 W and 4x3 Bh, 3x3 S
        W=np.zeros((N,4,4)); Bh=np.zeros((N,4,3)); S=np.zeros((N,3,3))
        W[:,0,0]=Ds[:,0]+Ds[:,1]; W[:,1,1]=Ds[:,2]+Ds[:,3]
        W[:,2,2]=(Ds[:,4:]*s.ax*s.ax).sum(1); W[:,2,3]=W[:,3,2]=(Ds[:,4:]*s.ax*s.ay).sum(1); W[:,3,3]=(Ds[:,4:]*s.ay*s.ay).sum(1)
        Bh[:,0,0]=Ds[:,1]; Bh[:,1,1]=Ds[:,2]-Ds[:,3]; Bh[:,2,2]=-(Ds[:,4:]*s.ax).sum(1); Bh[:,3,2]=-(Ds[:,4:]*s.ay).sum(1)
        S[:]=2*s.Qs; S[:,0,0]+=Ds[:,1]; S[:,1,1]+=Ds[:,2]+Ds[:,3]; S[:,2,2]+=Ds[:,4:].sum(1)
        Si=np.linalg.inv(S)
        Weff=W-np.einsum('tik,tkl,tjl->tij',Bh,Si,Bh)
        K=s.H+np.einsum('tia,tij,tjb->ab',s.Gb,Weff,s.Gb)+extra_reg*np.eye(nu)
        dbox=np.stack([Db[:,0]+Db[:,1],Db[:,2]+Db[:,3]],1).ravel()
        K=K+np.diag(dbox)
        rhs=ru-np.einsum('tia,tik,tkl,tl->a',s.Gb,Bh,Si,rsig)
        L=np.linalg.cholesky(K)
        du=np.linalg.solve(L.T,np.linalg.solve(L,rhs))
        do=np.einsum('tij,j->ti',s.Gb,du)
        dsig=np.einsum('tkl,tl->tk',Si,rsig-np.einsum('tik,ti->tk',Bh,do))
        return du,dsig
    def Arow(s,du,dsig):
        """A dv for all rows"""
        N=s.N
        do=np.einsum('tij,j->ti',s.Gb,du)
        rs=np.zeros((N,4+s.cnt))
        rs[:,0]=-do[:,0]; rs[:,1]=do[:,0]+dsig[:,0]; rs[:,2]=do[:,1]+dsig[:,1]; rs[:,3]=-do[:,1]+dsig[:,1]
        rs[:,4:]=s.ax*do[:,2:3]+s.ay*do[:,3:4]-dsig[:,2:3]
        U=du.reshape(N,2); rb=np.stack([U[:,0],-U[:,0],U[:,1],-U[:,1]],1)
        return rs,rb

def steplen(x,dx):
    neg=dx<0
    return min(1.0,(-x[neg]/dx[neg]).min()) if neg.any() else 1.0

def ipm(qp, mu_tol=1e-7, maxit=40, verbose=False):
    N=qp.N
    # init: weighted solve with D=1
    Ds=np.ones((N,4+qp.cnt)); Db=np.ones((N,4))
    u=np.zeros(2*N); sig=np.zeros((N,3))
    rs,rb=qp.resid(u,sig)
    # solve (H+A'A) v = -g - A' r0   (v from 0): minimise 1/2 v'Hv+g'v+1/2|Av-h|^2
    gu,gs=qp.grad_terms(rs,rb)
    du,dsig=qp.solve_weighted(Ds,Db,-qp.g-gu,-gs)
    u=u+du; sig=sig+dsig
    rs,rb=qp.resid(u,sig)
    ts=-rs; tb=-rb; ls=rs.copy(); lb=rb.copy()   # lam~ = -t~
    mn=min(ts.min(),tb.min())
    if mn<0: ts+=1-mn; tb+=1-mn
    mn=min(ls.min(),lb.min())
    if mn<0: ls+=1-mn; lb+=1-mn
    ts=np.maximum(ts,1e-8); tb=np.maximum(tb,1e-8); ls=np.maximum(ls,1e-8); lb=np.maximum(lb,1e-8)
    m=ts.size+tb.size
    for it in range(maxit):
        rs,rb=qp.resid(u,sig)
        rps=rs+ts; rpb=rb+tb
        gu,gs=qp.grad_terms(ls,lb)
        rdu=qp.H@u+qp.g+gu; rds=2*sig@qp.Qs.T+gs
        mu=((ls*ts).sum()+(lb*tb).sum())/m
        obj=abs(0.5*u@qp.H@u+qp.g@u)
        if verbose: print(it,'rd %.2e rp %.2e mu %.2e'%(max(np.abs(rdu).max(),np.abs(rds).max()),max(np.abs(rps).max(),np.abs(rpb).max()),mu))
        if mu<=mu_tol*(1+obj) and max(np.abs(rps).max(),np.abs(rpb).max())<=mu_tol*10 and max(np.abs(rdu).max(),np.abs(rds).max())<=mu_tol*(1+np.abs(qp.g).max()+max(ls.max(),lb.max()))*10:
            break
        Ds=ls/ts; Db=lb/tb
        def solve(rcs,rcb):
            ws=rcs/ts-Ds*rps; wb=rcb/tb-Db*rpb
            gu2,gs2=qp.grad_terms(ws,wb)
            du,dsig=qp.solve_weighted(Ds,Db,-rdu+gu2,-rds+gs2)
            As,Ab=qp.Arow(du,dsig)
            dts=-rps-As; dtb=-rpb-Ab
            dls=-(rcs+ls*dts)/ts; dlb=-(rcb+lb*dtb)/tb
            return du,dsig,dts,dtb,dls,dlb
        du,dsig,dts,dtb,dls,dlb=solve(ls*ts,lb*tb)
        ap=min(steplen(ts,dts),steplen(tb,dtb)); ad=min(steplen(ls,dls),steplen(lb,dlb))
        mu_aff=(((ls+ad*dls)*(ts+ap*dts)).sum()+((lb+ad*dlb)*(tb+ap*dtb)).sum())/m
        sg=(mu_aff/mu)**3
        du,dsig,dts,dtb,dls,dlb=solve(ls*ts+dts*dls-sg*mu, lb*tb+dtb*dlb-sg*mu)
        ap=min(1.0,0.99*min(steplen(ts,dts),steplen(tb,dtb))); ad=min(1.0,0.99*min(steplen(ls,dls),steplen(lb,dlb)))
        u=u+ap*du; sig=sig+ap*dsig; ts=ts+ap*dts; tb=tb+ap*dtb; ls=ls+ad*dls; lb=lb+ad*dlb
    return u,sig,ls,lb,ts,tb,it

def polish(qp, acts, actb, delta=1e-9, refine=4):
    """equality-constrained solve on active rows via regularised weighted normal equations + multiplier refinement"""
    N=qp.N
    Dv=1.0/delta
    Ds=np.where(acts,Dv,0.0); Db=np.where(actb,Dv,0.0)
    u=np.zeros(2*N); sig=np.zeros((N,3)); ls=np.zeros_like(Ds); lb=np.zeros_like(Db)
    for k in range(refine+1):
        rs,rb=qp.resid(u,sig)
        gu,gs=qp.grad_terms(ls+Ds*rs, lb+Db*rb)     # gradient of augmented Lagrangian at current point
        ru=-(qp.H@u+qp.g+gu); rsg=-(2*sig@qp.Qs.T+gs)
        du,dsig=qp.solve_weighted(Ds,Db,ru,rsg)
        u=u+du; sig=sig+dsig
        rs,rb=qp.resid(u,sig)
        ls=ls+Ds*rs; lb=lb+Db*rb
    return u,sig,ls,lb

def certify(qp,u,sig,ls,lb,acts,actb,tol=1e-9):
    rs,rb=qp.resid(u,sig)
    hs=1+np.abs(qp.hs); hb=1+np.abs(qp.hb)
    viols=(~acts)&(rs>tol*hs); violb=(~actb)&(rb>tol*hb)
    lsc=1+max(np.abs(ls).max(),np.abs(lb).max())
    negs=acts&(ls<-tol*lsc); negb=actb&(lb<-tol*lsc)
    ok=not(viols.any() or violb.any() or negs.any() or negb.any())
    return ok,viols,violb,negs,negb

def solve(qp, verbose=False, pdas_first=6, max_rounds=12):
    N=qp.N
    stats=dict(ipm=0,rounds=0)
    acts=np.zeros((N,4+qp.cnt),bool); actb=np.zeros((N,4),bool)
    def pdas(acts,actb,rounds):
        seen=set()
        for r in range(rounds):
            u,sig,ls,lb=polish(qp,acts,actb)
            stats['rounds']+=1
            ok,vs,vb,ns,nb=certify(qp,u,sig,ls,lb,acts,actb)
            if ok: return True,u,sig,ls,lb,acts,actb
            acts=(acts&~ns)|vs; actb=(actb&~nb)|vb
            key=acts.tobytes()+actb.tobytes()
            if key in seen: break
            seen.add(key)
        return False,u,sig,ls,lb,acts,actb
    ok,u,sig,ls,lb,acts,actb=pdas(acts,actb,pdas_first)
    mu_tol=1e-6
    while not ok and mu_tol>1e-13:
        u,sig,ls,lb,ts,tb,it=ipm(qp,mu_tol=mu_tol,verbose=verbose)
        stats['ipm']+=it
        acts=ls>ts; actb=lb>tb
        ok,u,sig,ls,lb,acts,actb=pdas(acts,actb,4)
        mu_tol*=1e-2
    return ok,u,sig,ls,lb,acts,actb,stats
