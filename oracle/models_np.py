"""Model log densities with analytic gradients (oracle; test infrastructure only).

The plugin contract follows ``make_stan_log_density`` (``viabel/vb.py:314-321``): a model
maps unconstrained ``x (n,d)`` to ``log p (n,)`` (Jacobian included, additive constants
free) and supplies ``grad log p (n,d)``.

Sources (citations into ``/root/reference``):
* funnel            – ``notebooks/funnel-distribution.ipynb:48-54`` (keeps norm.logpdf constants)
* normal mixture    – ``notebooks/normal-mixture.ipynb:43``
* Student-t regr.   – Stan program ``notebooks/robust-regression.ipynb:73-86``; ``~`` statements
                      drop additive constants; parameters unconstrained (no Jacobian)
* eight schools ncp – ``notebooks/eight_schools_ncp.stan:1-23``; data
                      ``notebooks/eight-schools.ipynb:173-174``; unconstrained order
                      (mu, log tau, theta_tilde[1..J]) with the +log tau Jacobian that
                      pystan's ``log_prob`` adds by default
* correlated Gauss  – not in the reference (BASELINE.json config 3); ``-1/2 (x-m)' P (x-m)``.
"""
import numpy as np

_LOG_2PI = np.log(2 * np.pi)

EIGHT_SCHOOLS_Y = np.array([28., 8., -3., 7., -1., 1., 18., 12.])
EIGHT_SCHOOLS_SIGMA = np.array([15., 10., 16., 11., 9., 11., 10., 18.])


class Model:
    """Callable ``model(x) -> (n,)`` like the reference's ``logdensity`` argument, plus
    ``logp_and_grad(x) -> ((n,), (n,d))``."""
    dim = None
    name = None

    def logp_and_grad(self, x):
        raise NotImplementedError

    def __call__(self, x):
        return self.logp_and_grad(np.atleast_2d(x))[0]


class Funnel(Model):
    name = 'funnel'
    dim = 2

    def __init__(self, log_sigma_stdev=1.35):
        self.a = float(log_sigma_stdev)

    def logp_and_grad(self, x):
        mu, ls = x[:, 0], x[:, 1]
        a = self.a
        e = np.exp(-2 * ls)
        # norm.logpdf(ls, 0, a) + norm.logpdf(mu, 0, exp(ls))
        lp = (-0.5 * _LOG_2PI - np.log(a) - 0.5 * (ls / a) ** 2) \
            + (-0.5 * _LOG_2PI - ls - 0.5 * mu * mu * e)
        g = np.stack([-mu * e, -ls / (a * a) - 1 + mu * mu * e], axis=1)
        return lp, g


class NormalMixture(Model):
    """log p = logaddexp(N(x;-2,1), N(x;2,1)) - log 2, one dimension."""
    name = 'mixture'
    dim = 1

    def __init__(self, loc=2.0):
        self.loc = float(loc)

    def logp_and_grad(self, x):
        t = x[:, 0]
        c = self.loc
        lp = np.logaddexp(-0.5 * (t + c) ** 2, -0.5 * (t - c) ** 2) - 0.5 * _LOG_2PI - np.log(2)
        g = -t + c * np.tanh(c * t)
        return lp, g[:, None]


class StudentTRegression(Model):
    """beta ~ N(0, 10); y ~ student_t(df, X beta, 1) — constants dropped."""
    name = 'regression'

    def __init__(self, X, y, df=40.0, prior_sd=10.0, row_block=1 << 15):
        self.X = np.ascontiguousarray(X, dtype=float)
        self.y = np.ascontiguousarray(y, dtype=float)
        self.df = float(df)
        self.prior_sd = float(prior_sd)
        self.dim = self.X.shape[1]
        self.row_block = int(row_block)

    def logp_and_grad(self, b):
        df = self.df
        n = b.shape[0]
        lp = -0.5 * np.sum(b * b, axis=1) / self.prior_sd ** 2
        g = -b / self.prior_sd ** 2
        N = self.X.shape[0]
        for s in range(0, N, self.row_block):          # never materialise (n, N)
            Xb = self.X[s:s + self.row_block]
            r = self.y[s:s + self.row_block][None, :] - b @ Xb.T        # (n, rows)
            lp = lp - 0.5 * (df + 1) * np.sum(np.log1p(r * r / df), axis=1)
            psi = (df + 1) * r / (df + r * r)
            g = g + psi @ Xb
        return lp, g


class EightSchoolsNCP(Model):
    name = 'eight_schools'

    def __init__(self, y=EIGHT_SCHOOLS_Y, sigma=EIGHT_SCHOOLS_SIGMA):
        self.y = np.asarray(y, dtype=float)
        self.sigma = np.asarray(sigma, dtype=float)
        self.dim = 2 + len(self.y)

    def logp_and_grad(self, u):
        mu, lt, th = u[:, 0], u[:, 1], u[:, 2:]
        tau = np.exp(lt)
        r = (self.y[None, :] - mu[:, None] - tau[:, None] * th) / self.sigma[None, :]
        ros = r / self.sigma[None, :]
        lp = -mu * mu / 50 - np.log1p(tau * tau / 25) + lt \
            - 0.5 * np.sum(th * th, axis=1) - 0.5 * np.sum(r * r, axis=1)
        g = np.empty_like(u)
        g[:, 0] = -mu / 25 + np.sum(ros, axis=1)
        g[:, 1] = 1 - 2 * tau * tau / (25 + tau * tau) + tau * np.sum(th * ros, axis=1)
        g[:, 2:] = -th + tau[:, None] * ros
        return lp, g


class CorrelatedGaussian(Model):
    """log p = -1/2 (x - m)' P (x - m), P symmetric positive definite (unnormalised)."""
    name = 'gauss'

    def __init__(self, mean, precision):
        self.m = np.asarray(mean, dtype=float)
        self.P = np.asarray(precision, dtype=float)
        self.dim = len(self.m)

    def logp_and_grad(self, x):
        dv = x - self.m[None, :]
        g = -(dv @ self.P)
        return 0.5 * np.sum(g * dv, axis=1), g
