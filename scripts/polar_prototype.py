import numpy as np, sys
sys.path.insert(0,'.')
SCHED=[(8.470329,-25.108079,18.629279),(4.182834,-3.108701,0.580607),(3.961856,-2.954063,0.562976),
       (3.286584,-2.464719,0.507358),(2.273748,-1.644659,0.416191),(1.888716,-1.265157,0.376519)]
MUON=(3.4445,-4.7750,2.0315)
def polar(L, verbose=False):
    d=L.shape[0]; I=np.eye(d); ng=0
    fro=np.linalg.norm(L); X=L/fro
    A=X.T@X; ng+=1
    s2=np.linalg.norm(A)*(1+1e-12); s=np.sqrt(s2)
    ps=s; err2=np.inf; hist=[]; status=1; fin=False
    for it in range(80):
        if it>0:
            A=X.T@X; ng+=1
            err2=np.linalg.norm(A-I)**2; hist.append(err2**.5)
            if err2<16*d*1e-28: status=0; break
        if it>=6 and (it-6)%5==0 and err2<=0.25: fin=True
        if err2<=0.0025: fin=True
        if not fin:
            a,b,c=SCHED[it if it<6 else 1+(it-6)%5]
            B=(b/ps**2)*A+(c/ps**4)*(A@A); ng+=1
            X=(a/ps)*X+(1/ps)*(X@B); ng+=1
        else:
            X=1.5*X-0.5*(X@A); ng+=1
            if err2<1e-14: status=0; break
        ps=1.0
    if verbose: print('its',it,'gemms',ng,'status',status,' '.join('%.1e'%h for h in hist))
    return X,ng
if __name__=='__main__':
    rs=np.random.RandomState(0)
    for name,kap,d in [('k1e2',1e2,300),('k1e4',1e4,300),('k1e6',1e6,200),('k1e9',1e9,200),('k1e12',1e12,128),('eye',1,64),('one',1,1),('two',10,2)]:
        Q,_=np.linalg.qr(rs.randn(d,d)); 
        L=(Q*np.geomspace(1/kap,1,d))@np.linalg.qr(rs.randn(d,d))[0]*3.7 if d>1 else np.array([[2.5]])
        X,ng=polar(L,True)
        U,sv,Vt=np.linalg.svd(L); Q0=U@Vt
        print(name,'cond %.1e'%np.linalg.cond(L),'Q err %.2e'%np.abs(X-Q0).max(), 'orth %.2e'%np.abs(X.T@X-np.eye(d)).max())
