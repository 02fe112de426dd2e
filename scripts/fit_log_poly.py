import numpy as np
from numpy.polynomial import chebyshev as C, polynomial as P
# log1p(f) = f - f^2/2 + f^3 Q(f)  on f in [2/3-1, 4/3-1];  fit Q (degree 6) by weighted least squares at Chebyshev nodes
a, b = 2/3 - 1, 4/3 - 1
n = 4000
k = np.arange(n)
t = np.cos(np.pi * (k + 0.5) / n)
f = 0.5 * (a + b) + 0.5 * (b - a) * t
f = f[np.abs(f) > 1e-6]
target = (np.log1p(f) - f + 0.5 * f * f) / f**3
best = None
for deg in (5, 6, 7):
    # minimise error of f^3 Q: weight by |f|^3
    w = np.abs(f) ** 3
    V = np.vander(f, deg + 1, increasing=True)
    coef, *_ = np.linalg.lstsq(V * w[:, None], target * w, rcond=None)
    # iterate (Lawson / remez-like reweighting)
    for it in range(60):
        err = (V @ coef - target) * np.abs(f) ** 3
        w = w * (1 + 4 * np.abs(err) / np.abs(err).max())
        coef, *_ = np.linalg.lstsq(V * w[:, None], target * w, rcond=None)
    c32 = coef.astype(np.float32)
    ff = np.linspace(a, b, 2000001)
    ff32 = ff.astype(np.float32)
    # emulate fp32 Horner
    q = np.full_like(ff32, c32[-1])
    for c in c32[-2::-1]:
        q = (q * ff32 + c).astype(np.float32)
    p = (q * ff32 - np.float32(0.5)).astype(np.float32)     # (Q f - 1/2)
    p = (p * ff32).astype(np.float32)                       # f (Q f - 1/2)
    res = (ff32 * p + ff32).astype(np.float32)              # f + f^2 (..)
    true = np.log1p(ff32.astype(np.float64))
    abserr = np.abs(res.astype(np.float64) - true)
    print(deg, 'max abs err', abserr.max(), 'mean signed', (res.astype(np.float64) - true).mean(), [float(x) for x in c32])
