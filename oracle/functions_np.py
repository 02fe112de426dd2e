"""numpy restatement of the reference's convergence diagnostics (viabel/functions.py) and of its two
R-hat-monitored optimisers (viabel/vb.py:392-712).  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

PINNED: against tests/golden/optimizers.npz, written by scripts/make_golden_optim.py from the reference's own
functions run in the build container (autograd.numpy stood in by numpy, which is all these functions use).
"""
import numpy as np

from .vb_np import learning_rate_schedule

__all__ = ['compute_R_hat', 'compute_R_hat_adaptive_numpy', 'compute_R_hat_halfway',
           'stochastic_iterate_averaging', 'rmsprop_IA_optimize_with_rhat', 'adam_IA_optimize_with_rhat']


def compute_R_hat(chains, warmup=500):
    """functions.py:8-31.  chains (n_chains, n_iters, K); every chain is cut in two halves (split R-hat).
    An odd number of kept iterations ends in the reference's own reshape error (its trim at :17-19 drops two
    rows, not one), restated here as the same ValueError."""
    jitter = 1e-8
    chains = np.asarray(chains)[:, warmup:, :]
    n_chains, n_iters, K = chains.shape
    if n_iters % 2 == 1:
        n_iters -= 1
        chains = chains[:, :n_iters - 1, :]                       # functions.py:19 (sic)
    n_iters //= 2
    psi = np.reshape(chains, (n_chains * 2, n_iters, K))
    m = n_chains * 2
    chain_mean = psi.mean(axis=1)
    grand_mean = chain_mean.mean(axis=0)
    s2 = ((psi - chain_mean[:, None, :]) ** 2).sum(axis=1) / (n_iters - 1)
    B = n_iters * ((chain_mean - grand_mean) ** 2).sum(axis=0) / (m - 1)
    W = np.nanmean(s2, axis=0) + jitter
    var_hat = (n_iters - 1) / n_iters + B / (n_iters * W)
    return var_hat, np.sqrt(var_hat)


def compute_R_hat_adaptive_numpy(chains, window_size=100):
    """functions.py:44-52: R-hat of every consecutive window of window_size iterations (n_iters must be a
    multiple of it: the reference reshapes)."""
    chains = np.asarray(chains)
    n_chains, n_iters, K = chains.shape
    n_windows = n_iters // window_size
    w = np.reshape(chains, [n_chains, n_windows, window_size, -1])
    return np.array([compute_R_hat(w[:, i], warmup=0)[1] for i in range(n_windows)])


def compute_R_hat_halfway(chains, interval=100, start=1000):
    """functions.py:54-65: R-hat of the second half of the growing prefixes chains[:, :start + (i+1) interval]."""
    chains = np.asarray(chains)
    n_chains, n_iters, K = chains.shape
    out = []
    for i in range(n_iters // interval):
        sub = chains[:, :start + (i + 1) * interval, :]
        out.append(compute_R_hat(sub, warmup=sub.shape[1] // 2)[1])
    return np.array(out)


def stochastic_iterate_averaging(estimate, start):
    """functions.py:68-77: running mean of estimate[start:] along the iterations.  The reference raises a str
    when start >= N, which Python turns into a TypeError."""
    estimate = np.asarray(estimate)
    N = estimate.shape[0]
    if N - start <= 0:
        raise TypeError('exceptions must derive from BaseException')
    lengths = np.reshape(np.arange(start, N) - start + 1, [-1, 1])
    iters = np.cumsum(estimate[start:, :], axis=0) / lengths
    return iters, iters[-1]


def _first_converged_window(rhat_windows, threshold, rhat_window, default):
    """vb.py:499-515: first window ee with it and the next one entirely below the threshold."""
    for ee in range(rhat_windows.shape[0] - 1):
        if (rhat_windows[ee] < threshold).all() and (rhat_windows[ee + 1] < threshold).all():
            return ee * rhat_window
    return default


def _optimize_with_rhat(kind, n_iters, objective_and_grad, init_param, K, has_log_norm, window, learning_rate,
                        epsilon, rhat_window, n_optimisers, r_mean_threshold, r_sigma_threshold, tail_avg_iters,
                        avg_grad_norm, learning_rate_end):
    from scipy import stats
    init_param = np.asarray(init_param, dtype=float)
    lam = init_param.copy()
    values, log_norms, chains = [], [], []
    alpha, beta1, beta2 = 0.9, 0.9, 0.999
    spread = 0.5 if kind == 'rmsprop' else 0.2                                   # vb.py:425 / :592
    for o in range(n_optimisers):
        history = []
        np.random.seed(seed=o)                                                   # vb.py:423 (global numpy stream)
        if o >= 1:
            lam = init_param + stats.norm.rvs(size=len(init_param)) * (o + 1) * spread
        for i, lr in enumerate(learning_rate_schedule(n_iters, learning_rate, learning_rate_end)):
            if has_log_norm:
                val, grad, log_norm = objective_and_grad(lam)
            else:
                val, grad = objective_and_grad(lam)
                log_norm = 0
            values.append(val)
            log_norms.append(log_norm)
            if kind == 'rmsprop':
                grad_norm = np.exp(log_norm) if has_log_norm else np.sum(grad ** 2, axis=0)
                if i == 0:
                    acc = grad_norm if avg_grad_norm else grad ** 2
                elif avg_grad_norm:
                    acc = grad_norm * alpha + (1. - alpha) * grad_norm            # vb.py:452 (sic)
                else:
                    acc = acc * alpha + (1. - alpha) * grad ** 2
                step = lr * grad / np.sqrt(epsilon + acc)                        # vb.py:461-462
            else:
                if i == 0:
                    v, m = 0.9 * grad ** 2, 0.9 * grad                           # vb.py:617-618
                else:
                    v = v * beta2 + (1. - beta2) * grad ** 2
                    m = m * beta1 + (1. - beta1) * grad
                m_hat = m / (1 - np.power(beta1, i + 2))
                v_hat = v / (1 - np.power(beta2, i + 2))
                step = lr * m_hat / np.sqrt(epsilon + v_hat)                     # vb.py:629
            history.append(lam.copy())
            lam = lam - step
            if len(history) > 100 * window:
                history.pop(0)
        chains.append(np.array(history))
    chains = np.stack(chains, axis=0)
    rhats = compute_R_hat_adaptive_numpy(chains, window_size=rhat_window)
    halfway = compute_R_hat_halfway(chains, interval=100, start=200)
    start_m = _first_converged_window(rhats[:, :K], r_mean_threshold, rhat_window, n_iters - tail_avg_iters)
    start_s = _first_converged_window(rhats[:, K:], r_sigma_threshold, rhat_window, n_iters - tail_avg_iters)
    avg_means, avg_sigmas = [], []
    for o in range(n_optimisers):
        avg_means.append(stochastic_iterate_averaging(chains[o, :, :K], start_m)[0])
        avg_sigmas.append(stochastic_iterate_averaging(chains[o, :, K:], start_s)[0])
    log = {'start_avg_mean_iters': start_m, 'start_avg_sigma_iters': start_s,
           'r_hat_mean': rhats[:, :K], 'r_hat_sigma': rhats[:, K:],
           'r_hat_mean_halfway': halfway[:, :K], 'r_hat_sigma_halfway': halfway[:, K:]}
    return lam, chains, avg_means, avg_sigmas, np.array(values), np.array(log_norms), log


def rmsprop_IA_optimize_with_rhat(n_iters, objective_and_grad, init_param, K, has_log_norm=False, window=500,
                                  learning_rate=.01, epsilon=.000001, rhat_window=500, averaging=True,
                                  n_optimisers=1, r_mean_threshold=1.15, r_sigma_threshold=1.20,
                                  tail_avg_iters=2000, avg_grad_norm=False, learning_rate_end=None):
    """vb.py:392-548 without the progress bar and the unused windowed-gradient bookkeeping (:456-459)."""
    return _optimize_with_rhat('rmsprop', n_iters, objective_and_grad, init_param, K, has_log_norm, window,
                               learning_rate, epsilon, rhat_window, n_optimisers, r_mean_threshold,
                               r_sigma_threshold, tail_avg_iters, avg_grad_norm, learning_rate_end)


def adam_IA_optimize_with_rhat(n_iters, objective_and_grad, init_param, K, has_log_norm=False, window=500,
                               learning_rate=.01, epsilon=.000001, rhat_window=500, averaging=True,
                               n_optimisers=1, r_mean_threshold=1.15, r_sigma_threshold=1.20,
                               tail_avg_iters=2000, learning_rate_end=None):
    """vb.py:554-712 likewise (the in-loop `averaged_variational_param` of :631-634 is never returned)."""
    return _optimize_with_rhat('adam', n_iters, objective_and_grad, init_param, K, has_log_norm, window,
                               learning_rate, epsilon, rhat_window, n_optimisers, r_mean_threshold,
                               r_sigma_threshold, tail_avg_iters, False, learning_rate_end)
