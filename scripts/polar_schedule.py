import numpy as np
from scipy.optimize import linprog
np.set_printoptions(precision=6, linewidth=150)

def opt_quintic(l, u, ngrid=4000):
    # minimise max |1 - (a x + b x^3 + c x^5)| over [l, u] (log-spaced + linear grid)
    x = np.unique(np.concatenate([np.geomspace(l, u, ngrid), np.linspace(l, u, ngrid)]))
    P = np.stack([x, x**3, x**5], 1)
    # variables a,b,c,t ; constraints  P@coef - 1 <= t ; 1 - P@coef <= t
    A = np.block([[P, -np.ones((len(x),1))],[-P, -np.ones((len(x),1))]])
    bvec = np.concatenate([np.ones(len(x)), -np.ones(len(x))])
    res = linprog([0,0,0,1], A_ub=A, b_ub=bvec, bounds=[(None,None)]*4, method='highs')
    a,b,c,t = res.x
    return (a,b,c), t

def schedule(l0, n, safety=1.0):
    l, u = l0, 1.0
    out = []
    for k in range(n):
        (a,b,c), t = opt_quintic(l, u)
        a,b,c = a/safety, b/safety**3, c/safety**5   # p(x/safety)
        out.append((a,b,c))
        l, u = (1 - t), (1 + t)
        print(k, 'coef %.6f %.6f %.6f' % (a,b,c), 'err %.3e' % t)
        if t < 1e-16: break
    return out

if __name__ == '__main__':
    import sys
    l0 = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-3
    schedule(l0, 10)
