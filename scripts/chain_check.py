import sys, numpy as np
sys.path.insert(0,'.')
def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)
def _gauss(d, seed=4):
    rs = np.random.RandomState(seed)
    A = rs.randn(d, d)
    Sp = A @ A.T / d + np.eye(d)
    mp = np.random.RandomState(seed + 1).randn(d)
    return mp, np.linalg.inv(Sp)
from oracle import vb_np, models_np as mm
from viabel_b200 import vb, models
for span in [1.5,3.,4.5,5.0]:
    d,n=40,2000
    mp,Pp=_gauss(d)
    fam, ofam = vb.t_variational_family(d, 40), vb_np.t_variational_family(d, 40)
    rs=np.random.RandomState(8); lam=0.05*rs.randn(fam.var_param_dim)
    r,c=np.tril_indices(d); dp=d+np.flatnonzero(r==c); lam[dp]=np.linspace(-span,0.5,d)[rs.permutation(d)]
    _,L,_=ofam.impl.fold(lam)
    base=ofam.impl.draw_base(np.random.RandomState(5),n)
    v,g=vb.black_box_chivi(2.0,fam,models.CorrelatedGaussian(mp,Pp),n,group=False)(lam,base_draws=base)
    ov,og=vb_np.chivi_value_and_grad(ofam.impl,mm.CorrelatedGaussian(mp,Pp),2.0,lam,base)
    print('span',span,'cond %.2e'%np.linalg.cond(L),'info',fam.impl.chain_info(),'v rel %.2e'%abs(v/ov-1),'g rel %.2e'%rel(g,og))
