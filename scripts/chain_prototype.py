import numpy as np, sys
sys.path.insert(0,'/root/repo')
import bench_workloads as W
from oracle import vb_np
np.set_printoptions(precision=3)

def chain(L, G, verbose=False):
    d=L.shape[0]; I=np.eye(d); ng=0
    fro=np.linalg.norm(L)
    X=L/fro
    # quintic lift (fixed 6) then Newton-Schulz until converged
    a,b,c=3.4445,-4.7750,2.0315
    for it in range(6):
        A=X.T@X; ng+=1
        err=np.linalg.norm(A-I)
        if err<0.5: break
        B=b*A+c*(A@A); ng+=1
        X=a*X+X@B; ng+=1
    for it in range(40):
        A=X.T@X; ng+=1
        err=np.linalg.norm(A-I)
        if err<1e-14*np.sqrt(d)*4: break
        X=1.5*X-0.5*(X@A); ng+=1
    polar_its=it
    S=L@X.T; ng+=1
    S=0.5*(S+S.T)
    # Sylvester S Y + Y S = G
    mu=fro/np.sqrt(d)
    M=S+mu*I
    n1=np.abs(M).sum(0).max()
    Xi=I/n1
    for it in range(40):
        R=M@Xi; ng+=1
        err=np.linalg.norm(R-I)
        if err<1e-14*np.sqrt(d)*4: break
        Xi=2*Xi-Xi@R; ng+=1
    inv_its=it
    Aq=S@Xi-mu*Xi; ng+=1
    T=Xi@G; ng+=1
    Y=2*mu*(T@Xi); ng+=1
    for it in range(40):
        nA=np.linalg.norm(Aq)
        if nA<1e-17: break
        T=Aq@Y; ng+=1
        Y=Y+T@Aq; ng+=1
        Aq=Aq@Aq; ng+=1
    smith_its=it
    if verbose: print('polar its',polar_its,'inv its',inv_its,'smith its',smith_its,'GEMMs',ng)
    return S,Y,ng

def ref(L,G):
    Sigma=L@L.T
    lam,V=np.linalg.eigh(Sigma); r=np.sqrt(lam)
    S=(V*r)@V.T
    Y=V@((V.T@G@V)/(r[:,None]+r[None,:]))@V.T
    return S,Y

for name in ['c3','identity','illcond','d24']:
    if name=='c3':
        mp,Pp,lam=W.c3_problem(512); fam=vb_np.t_variational_family(512,40); mu_,L,Sig=fam.impl.fold(lam)
    elif name=='identity':
        L=np.eye(512)
    elif name=='illcond':
        rs=np.random.RandomState(0); Q,_=np.linalg.qr(rs.randn(300,300)); L=np.linalg.cholesky((Q*np.logspace(-6,1,300))@Q.T)
    else:
        rs=np.random.RandomState(1); L=np.tril(0.3*rs.randn(24,24)); L[np.diag_indices(24)]=np.exp(np.diag(L))
    d=L.shape[0]
    G=np.random.RandomState(3).randn(d,d)
    S,Y,ng=chain(L,G,True)
    S0,Y0=ref(L,G)
    print(name,'cond(L)=%.1e'%np.linalg.cond(L),'S err %.2e'%(np.abs(S-S0).max()/np.abs(S0).max()),'Y err %.2e'%(np.abs(Y-Y0).max()/np.abs(Y0).max()),
          'resid %.2e'%(np.abs(S@Y+Y@S-G).max()/np.abs(G).max()), 'ref resid %.2e'%(np.abs(S0@Y0+Y0@S0-G).max()/np.abs(G).max()))
