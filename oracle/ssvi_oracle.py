"""CPU oracle: a numpy restatement of SSVI-TF's per-row stochastic variational update.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package
(ssvi_tf_b200/) never does and fails loudly when its CUDA library is missing.

PARITY PINNING: the reference holds no golden vectors or known-answer tests for this
path.  This restatement is pinned against OUTPUTS OF THE
REFERENCE ITSELF: oracle/gen_golden.py runs the unmodified reference's
update_natural_param_batch under the noise tape of oracle/ref_shim.py in the build
container and commits inputs+outputs under tests/golden/*.npz;
tests/test_oracle_golden.py checks every function here against those files.

Each function cites the reference lines it restates (paths relative to /root/reference).
The noise contract (standard variates supplied by the caller, transform x = m + L z with L
the lower Cholesky factor) is described in oracle/ref_shim.py and DESIGN.md.

State dict (same fields the reference mutates, SSVI/SSVI_TF_interface.py:38-98):
  params[d]        (n_d, D+1, D) full: row 0 mean, rows 1..D covariance; diag: (n_d, 2, D)
  ada_mean[d]      (D, n_d)   ada_acc_grad           (init ones, :83)
  ada_cov[d]       (D, n_d)   ada_acc_grad_cov       (diag only, init zeros, :41)
  norm_changes[d]  (n_d, 2)   grad_norms[d] (n_d, 2)
  pSigma (N,)  pmu (D,)  time_step [N]  w_sigma  w_ada_grad  w_tau
"""
import math

import numpy as np

LIK_NORMAL, LIK_BERNOULLI, LIK_POISSON = "normal", "bernoulli", "poisson"
DATATYPE_TO_LIK = {"real": LIK_NORMAL, "binary": LIK_BERNOULLI, "count": LIK_POISSON}


def make_config(model, likelihood, D, diag=False, k1=64, k2=64, batch_size=128,
                mean_update="S", cov_update="N", eta=1.0, cov_eta=1.0, sigma_eta=1.0, quadrature=0):
    """Constructor kwargs + hidden constants of SSVI_TF.__init__ (SSVI_TF_interface.py:21-98)."""
    assert model in ("d", "simple", "robust")
    assert likelihood in (LIK_NORMAL, LIK_BERNOULLI, LIK_POISSON)
    return dict(model=model, likelihood=likelihood, D=int(D), diag=bool(diag), k1=int(k1), k2=int(k2),
                batch_size=int(batch_size), mean_update=mean_update, cov_update=cov_update,
                eta=float(eta), cov_eta=float(cov_eta), sigma_eta=float(sigma_eta),
                poisson_eta=0.1,            # :87
                min_eigenval=0.01,          # :48
                epsilon=1e-4,               # :80
                likelihood_param=1.0 if likelihood == LIK_NORMAL else 0.0,   # :62, :67, :76
                noise_added=(model != "d"),  # SSVI_TF_simple.py:18, SSVI_TF_robust.py:19
                # SSVI_TF_robust(quadrature=True): Gauss-Hermite degree (the reference fixes 30, :23); 0 = Monte Carlo
                quadrature=int(quadrature) if model == "robust" else 0)


def new_state(cfg, dims, rng=None, mean0=None, cov0=None, randstart=True):
    """Posterior_{Full,Diag}_Covariance.initialize_params (Model/TF_Models.py:53-62, 77-86)
    plus the trackers of SSVI_TF.__init__.  rng: numpy Generator used for the random start."""
    D = cfg["D"]
    mean0 = np.ones(D) if mean0 is None else np.asarray(mean0, float)
    cov0 = np.eye(D) if cov0 is None else np.asarray(cov0, float)
    rng = np.random.default_rng(0) if rng is None else rng
    Lc = np.linalg.cholesky(cov0)
    params = []
    for n in dims:
        if cfg["diag"]:
            p = np.zeros((n, 2, D))
            p[:, 1, :] = np.diag(cov0)
        else:
            p = np.zeros((n, D + 1, D))
            p[:, 1:, :] = cov0
        p[:, 0, :] = mean0 + (rng.standard_normal((n, D)) @ Lc.T if randstart else 0.0)
        params.append(p)
    return dict(params=params,
                ada_mean=[np.ones((D, n)) for n in dims],
                ada_cov=[np.zeros((D, n)) for n in dims],
                norm_changes=[np.zeros((n, 2)) for n in dims],
                grad_norms=[np.zeros((n, 2)) for n in dims],
                pSigma=np.ones(len(dims)), pmu=np.ones(D), time_step=[0] * len(dims),
                w_sigma=1.0, w_ada_grad=0.0, w_tau=1.0, fixed_covariance=False)


def copy_state(s):
    out = dict(s)
    for k in ("params", "ada_mean", "ada_cov", "norm_changes", "grad_norms"):
        out[k] = [a.copy() for a in s[k]]
    out["pSigma"] = s["pSigma"].copy()
    out["pmu"] = s["pmu"].copy()
    out["time_step"] = list(s["time_step"])
    return out


# --------------------------------------------------------------------------------------
# Link functions (Probability/ProbFun.py)
# --------------------------------------------------------------------------------------
def sigmoid(x, epsilon=0.0):
    """ProbFun.py:22-28: 1/(1+exp(-x)), nan_to_num, clipped to [epsilon, 1-epsilon]."""
    with np.errstate(all="ignore"):
        res = 1.0 / (1.0 + np.exp(-x))
    res = np.nan_to_num(res)
    return np.minimum(np.maximum(res, epsilon), 1.0 - epsilon)


def poisson_link(f, epsilon=1e-8):
    """ProbFun.py:52-55: log(1+exp(f)) (NOT log1p), nan_to_num, floored at 1e-8."""
    with np.errstate(all="ignore"):
        res = np.log(1.0 + np.exp(f))
    res = np.nan_to_num(res)
    return np.maximum(res, epsilon)


_SQRT_2PI = math.sqrt(2.0 * math.pi)


# --------------------------------------------------------------------------------------
# Likelihood terms.  y scalar or broadcastable, f array, s = likelihood_param
# --------------------------------------------------------------------------------------
def fst_derivative_log_pdf(lik, y, f, s):
    if lik == LIK_NORMAL:            # normal.py:32-33
        return (y - f) / (s * s)
    if lik == LIK_BERNOULLI:         # bernoulli.py:40-47
        return y * sigmoid(-(y * f))
    sg = sigmoid(f)                  # poisson.py:70-77
    A = poisson_link(f)
    with np.errstate(all="ignore"):
        return np.nan_to_num(sg * (y / A - 1.0))


def snd_derivative_log_pdf(lik, y, f, s):
    if lik == LIK_NORMAL:            # normal.py:35-36
        return -1.0 / (s * s) * np.ones_like(f)
    if lik == LIK_BERNOULLI:         # bernoulli.py:49-51
        return -sigmoid(y * f) * sigmoid(-(y * f))
    sg = sigmoid(f)                  # poisson.py:79-87
    A = poisson_link(f)
    with np.errstate(all="ignore"):
        return np.nan_to_num(sg * (1.0 - sg) * (y / A - 1.0) - y * np.square(sg) / np.square(A))


def _poisson_pdf(y, m):
    """poisson.py:19-27: A^y / (1+e^m) / y!, floored at 1e-8 (epsilon)."""
    A = poisson_link(m)
    with np.errstate(all="ignore"):
        res = np.power(A, y) / (1.0 + np.exp(m)) / math.factorial(int(y))
    return np.maximum(res, 1e-8)


def pdf(lik, y, f, s):
    if lik == LIK_NORMAL:            # normal.py:16-18 (scipy.stats.norm.pdf floored at 1e-5)
        r = (y - f) / s
        return np.maximum(np.exp(-0.5 * r * r) / (_SQRT_2PI * s), 1e-5)
    if lik == LIK_BERNOULLI:         # bernoulli.py:16-21
        return sigmoid(y * f, 1e-16)
    return _poisson_pdf(y, f)


def fst_derivative_pdf(lik, y, f, s):
    if lik == LIK_NORMAL:            # normal.py:20-23
        return pdf(lik, y, f, s) * ((y - f) / (s * s))
    if lik == LIK_BERNOULLI:         # bernoulli.py:23-28
        sg = sigmoid(y * f, 1e-16)
        return (y * sg) * (1.0 - sg)
    A = poisson_link(f)              # poisson.py:29-39
    return _poisson_pdf(y, f) * (sigmoid(f) * (y / A - 1.0))


def snd_derivative_pdf(lik, y, f, s):
    if lik == LIK_NORMAL:            # normal.py:25-27 (returns pdf*(((y-m)/s)^2 - 1/s); quirk Q5)
        return pdf(lik, y, f, s) * (np.square((y - f) / s) - 1.0 / s)
    if lik == LIK_BERNOULLI:         # bernoulli.py:30-36
        sg = sigmoid(y * f, 1e-16)
        return sg * (1.0 - sg) * (1.0 - 2.0 * sg)
    # poisson.py:41-57; the pdf is evaluated at poisson_link(m): link applied twice (quirk Q7)
    p2 = _poisson_pdf(y, poisson_link(f))
    sg = sigmoid(f)
    A = poisson_link(f)
    pmf_prime = fst_derivative_pdf(lik, y, f, s)
    temp1 = pmf_prime * (sg * (y / A - 1.0))
    temp = -(y / np.square(A)) * np.square(sg) + (y / A - 1.0) * (sg * (1.0 - sg))
    return temp1 + p2 * temp


# --------------------------------------------------------------------------------------
# Posterior access (Model/TF_Models.py:64-69, 88-93)
# --------------------------------------------------------------------------------------
def get_vector_distribution(state, cfg, dim, i):
    p = state["params"][dim]
    if cfg["diag"]:
        return p[i, 0, :], p[i, 1, :]
    return p[i, 0, :], p[i, 1:, :]


def _chol(cfg, S):
    """Sample transform of the noise contract: L with x = m + L z."""
    if cfg["diag"]:
        return np.diag(np.sqrt(S))
    return np.linalg.cholesky(S)


# --------------------------------------------------------------------------------------
# Stage 1+2: sampling (SSVI_TF_interface.py:732-750, :251-257; SSVI_TF_robust.py:42-56)
# --------------------------------------------------------------------------------------
def sample_vjs_batch(state, cfg, dim, coords, z_other):
    """vjs[b,k,:] = prod over other modes (ascending) of x = m_j + L_j z_other[b,jj,k,:]."""
    B, N = coords.shape
    k1, D = cfg["k1"], cfg["D"]
    vjs = np.ones((B, k1, D))
    otherdims = [d for d in range(N) if d != dim]
    for b in range(B):
        for jj, d in enumerate(otherdims):
            mj, Sj = get_vector_distribution(state, cfg, d, coords[b, d])
            vjs[b] *= mj + z_other[b, jj] @ _chol(cfg, Sj).T
    return vjs


def estimate_di_Di_si(state, cfg, dim, m, S, coords, ys, tape):
    """di, Di, si for one row minibatch.

    d / simple: estimate_di_Di_si_batch_full_samplings (SSVI_TF_interface.py:229-264) followed by
      SSVI_TF_d.approximate_di_Di_si_without_second_layer_samplings (SSVI_TF_d.py:31-57) or
      approximate_di_Di_si_with_second_layer_samplings (:562-592) with
      SSVI_TF_simple.estimate_expected_derivative_batch (SSVI_TF_simple.py:23-40).
    robust: SSVI_TF_robust.estimate_di_Di_si_batch (:28-65), :67-145, :147-169.
    """
    lik, s = cfg["likelihood"], cfg["likelihood_param"]
    k1, k2, D = cfg["k1"], cfg["k2"], cfg["D"]
    B = coords.shape[0]
    vjs = sample_vjs_batch(state, cfg, dim, coords, tape.z_other)            # (B,k1,D)
    uis = m + tape.z_own @ _chol(cfg, S).T                                   # (B,k1,D)
    mean_batch = np.sum(uis * vjs, axis=2)                                   # (B,k1)
    ys = np.asarray(ys, float)

    if cfg["model"] == "d":
        # Q3: np.average(axis=0) of a (k1,) vector is a scalar, broadcast over k1
        fst = np.empty((B, k1))
        snd = np.empty((B, k1))
        for b in range(B):
            fst[b] = np.average(fst_derivative_log_pdf(lik, ys[b], mean_batch[b], s))
            snd[b] = np.average(snd_derivative_log_pdf(lik, ys[b], mean_batch[b], s))
        a_w = fst / k1                    # di weight per (b,k)
        D_w = 0.5 * snd / k1              # Di weight per (b,k)
        si = None
    elif cfg["model"] == "simple":
        w_sigma = state["w_sigma"]
        ws = (w_sigma ** 2) * tape.r_w                                       # :260 rayleigh(scale=w_sigma^2)
        fst = np.empty((B, k1))
        snd = np.empty((B, k1))
        for b in range(B):
            fs = mean_batch[b] + (0.0 + ws[b]) * tape.z_f[b]                 # Q2: scale = var + ws
            fst[b] = np.average(fst_derivative_log_pdf(lik, ys[b], fs, s), axis=0)
            snd[b] = np.average(snd_derivative_log_pdf(lik, ys[b], fs, s), axis=0)
        si = np.mean(snd * ws / (8.0 * w_sigma ** 2))                        # SSVI_TF_simple.py:38-39 (a MEAN)
        a_w = fst / k1
        D_w = 0.5 * snd / k1
    else:
        w_sigma = state["w_sigma"]
        ws = (w_sigma ** 2) * tape.r_w                                       # SSVI_TF_robust.py:53
        phi = np.empty((B, k1))
        phi1 = np.empty((B, k1))
        phi2 = np.empty((B, k1))
        if cfg.get("quadrature", 0):
            # estimate_expected_derivatives_pdf_batch_quadrature / estimate_expected_pdf_quadrature (:175-214):
            # fs = m + sqrt(2 w^2) xs;  E = sum(weights * g(fs)) / sqrt(pi); no second-layer draws are consumed
            xs, wts = np.polynomial.hermite.hermgauss(int(cfg["quadrature"]))            # :23-24
            for b in range(B):
                fs = mean_batch[b][None, :] + np.sqrt(2 * np.square(ws[b]))[None, :] * xs[:, None]   # (degree, k1)
                phi[b] = np.sum(wts[:, None] * pdf(lik, ys[b], fs, s), axis=0) / np.sqrt(np.pi)
                phi1[b] = np.sum(wts[:, None] * fst_derivative_pdf(lik, ys[b], fs, s), axis=0) / np.sqrt(np.pi)
                phi2[b] = np.sum(wts[:, None] * snd_derivative_pdf(lik, ys[b], fs, s), axis=0) / np.sqrt(np.pi)
        for b in range(B if not cfg.get("quadrature", 0) else 0):
            fs = mean_batch[b] + ws[b] * tape.z_f[b]                         # :163
            phi[b] = np.mean(pdf(lik, ys[b], fs, s), axis=0)
            phi1[b] = np.mean(fst_derivative_pdf(lik, ys[b], fs, s), axis=0)
            phi2[b] = np.mean(snd_derivative_pdf(lik, ys[b], fs, s), axis=0)
        assert np.all(np.isfinite(phi)) and np.all(np.isfinite(phi1)) and np.all(np.isfinite(phi2))  # :87-92
        assert not np.any(phi == 0)                                          # :106
        a_w = (phi1 / phi) / k1                                              # :110-111
        si = np.sum(np.mean(ws * phi2 / phi, axis=1) / (8.0 * w_sigma ** 2))  # :113, :137 (a SUM over b)
        # :115-133  0.5/(k1 p) * (p2 v v^T - (1/p)(p2 v)(p2 v)^T)  (Q6: p2 where p1 is expected)
        D_w = 0.5 / (k1 * phi) * (phi2 - phi2 * phi2 / phi)

    di = np.einsum("bk,bkd->d", a_w, vjs)
    if cfg["diag"]:
        Di = np.einsum("bk,bkd->d", D_w, vjs * vjs)
    else:
        Di = np.einsum("bk,bkd,bke->de", D_w, vjs, vjs)
    return di, Di, si


# --------------------------------------------------------------------------------------
# Stage 4: row update
# --------------------------------------------------------------------------------------
def _adagrad_step(acc_col, grad, eta):
    """compute_stepsize_mean_param (:694-698) incl. the aliasing quirk Q1:
    acc_grad is a VIEW, incremented in place before the division, so
    step = eta / sqrt(acc_old + 2 g^2) while the stored accumulator is acc_old + g^2."""
    g2 = np.square(grad)
    acc_col += g2
    return eta / np.sqrt(acc_col + g2)


def update_cov_param(state, cfg, dim, i, m, S, di, Di):
    """update_cov_param (:616-637) / update_cov_param_diag (:661-681)."""
    D = cfg["D"]
    sigma = state["pSigma"][dim]
    if cfg["diag"]:
        pinv = np.full((D,), 1.0 / sigma)
        if cfg["cov_update"] == "S":          # Q8: L_next computed then discarded
            L = np.sqrt(S)
            covGrad = 1.0 / L - L * pinv + 2.0 * L * Di
            _adagrad_step(state["ada_cov"][dim][:, i], covGrad, 1.0)
            S_next = np.square(L)
        elif cfg["cov_update"] == "N":
            covGrad = pinv - 2.0 * Di
            eta = cfg["poisson_eta"] if cfg["likelihood"] == LIK_POISSON else cfg["eta"]   # :710-713
            step = _adagrad_step(state["ada_cov"][dim][:, i], covGrad, eta)
            S_next = 1.0 / ((1.0 - step) * (1.0 / S) + step * covGrad)
        else:
            raise Exception("Unidentified update formula for covariance param")
        return np.maximum(S_next, cfg["epsilon"]), covGrad
    pinv = np.eye(D) / sigma
    if cfg["cov_update"] == "S":              # Q8
        L = np.linalg.cholesky(S)
        covGrad = np.triu(np.linalg.inv(L * np.eye(D)) - np.inner(L, pinv) + 2.0 * np.dot(L, Di))
        S_next = L @ L.T
    elif cfg["cov_update"] == "N":
        covGrad = pinv - 2.0 * Di
        step = cfg["cov_eta"] / (state["time_step"][dim] + 1)                # :701-702
        S_next = np.linalg.inv((1.0 - step) * np.linalg.inv(S) + step * covGrad)
    else:
        raise Exception("Unidentified update formula for covariance param")
    w, v = np.linalg.eigh(S_next)                                            # :634-636
    w = np.maximum(np.real(w), cfg["min_eigenval"])
    return v @ np.diag(w) @ v.T, covGrad


def update_mean_param(state, cfg, dim, i, m, S, di, Di):
    """update_mean_param (:595-613) / update_mean_param_diag (:640-658)."""
    sigma = state["pSigma"][dim]
    pmu = state["pmu"]
    acc = state["ada_mean"][dim][:, i]
    if cfg["mean_update"] == "S":
        grad = (pmu - m) / sigma + di
        step = _adagrad_step(acc, grad, cfg["eta"])
        return m + step * grad, grad
    if cfg["mean_update"] == "N":
        if cfg["diag"]:
            C = m / S
            grad = pmu / sigma + di - 2.0 * Di * m
            step = _adagrad_step(acc, grad, cfg["eta"])
            return S * ((1.0 - step) * C + step * grad), grad
        C = np.linalg.inv(S) @ m
        grad = pmu / sigma + di - 2.0 * (Di @ m)
        step = _adagrad_step(acc, grad, cfg["eta"])
        return S @ ((1.0 - step) * C + step * grad), grad
    raise Exception("Unidentified update formula for covariance param")


def update_sigma_param(state, cfg, si, scale):
    """update_sigma_param (:684-691), robust override (SSVI_TF_robust.py:216-232),
    compute_stepsize_sigma_param (:716-719).  Mutates w_sigma / w_ada_grad."""
    si = si * scale
    w_grad = -1.0 / (2.0 * state["w_tau"] ** 2) + si
    state["w_ada_grad"] += w_grad * w_grad
    w_step = cfg["sigma_eta"] / state["w_ada_grad"]
    update = (1.0 - w_step) * (-0.5 / state["w_sigma"] ** 2) + w_step * w_grad
    with np.errstate(all="ignore"):
        nxt = float(np.sqrt(-0.5 / update))
    if cfg["model"] == "robust" and math.isnan(nxt):
        return
    state["w_sigma"] = nxt


def row_accumulate(state, cfg, dim, row, coords_row, ys_row, tape):
    """Stages 1-3 for one row: returns (di, Di, si, scale, B) BEFORE the scale multiply.
    coords_row/ys_row are ALL observed entries of the row in index order
    (Tensor.find_observed_ui, Tensor/Tensor.py:251-258); the minibatch is tape.sub_idx when the
    row holds more than batch_size entries (:170-174)."""
    n_obs = len(ys_row)
    if n_obs > cfg["batch_size"]:
        sub = np.asarray(tape.sub_idx)
        coords, ys = coords_row[sub], ys_row[sub]
    else:
        coords, ys = coords_row, ys_row
    m, S = get_vector_distribution(state, cfg, dim, row)
    di, Di, si = estimate_di_Di_si(state, cfg, dim, m, S, coords, ys, tape)
    return di, Di, si, n_obs / len(ys), len(ys)


def row_update(state, cfg, dim, row, di, Di, scale):
    """Stage 4 for one row (:181-219 without the sigma update): mutates params, AdaGrad state
    and the trackers of row `row`."""
    m, S = get_vector_distribution(state, cfg, dim, row)
    m, S = m.copy(), S.copy()
    di = di * scale
    Di = Di * scale
    assert np.all(np.isfinite(Di))                                           # :185-188
    S_next, S_grad = update_cov_param(state, cfg, dim, row, m, S, di, Di)
    m_next, m_grad = update_mean_param(state, cfg, dim, row, m, S, di, Di)
    state["grad_norms"][dim][row] = [np.linalg.norm(m_grad), np.linalg.norm(S_grad)]      # :199
    assert np.all(np.isfinite(S_next))                                       # :202-204
    if state.get("fixed_covariance", False):                                 # :213-214
        S_next = S
    state["norm_changes"][dim][row] = [np.linalg.norm(m_next - m), np.linalg.norm(S_next - S)]  # :722-729
    p = state["params"][dim]                                                 # :219
    p[row, 0, :] = m_next
    if cfg["diag"]:
        p[row, 1, :] = S_next
    else:
        p[row, 1:, :] = S_next


def row_step(state, cfg, dim, row, coords_row, ys_row, tape):
    """update_natural_param_batch(dim, row) (:163-219).  Returns False for an empty row (:166-168)."""
    if len(ys_row) == 0:
        return False
    di, Di, si, scale, _ = row_accumulate(state, cfg, dim, row, coords_row, ys_row, tape)
    row_update(state, cfg, dim, row, di, Di, scale)
    if cfg["noise_added"]:                                                   # :207-210
        update_sigma_param(state, cfg, si, scale)
    return True


def rows_step(state, cfg, dim, rows, row_entries, tapes):
    """The batched equivalence class of SURVEY.md section 8b: rows of ONE mode, all evaluated from
    the same snapshot (w_sigma, pSigma, time_step frozen), then the scalar w_sigma recurrence
    replayed in row order.  row_entries[r] = (coords_row, ys_row); tapes[r] = NoiseTape-like."""
    pending = []
    for r, row in enumerate(rows):
        coords_row, ys_row = row_entries[r]
        if len(ys_row) == 0:
            continue
        di, Di, si, scale, _ = row_accumulate(state, cfg, dim, row, coords_row, ys_row, tapes[r])
        pending.append((row, di, Di, si, scale))
    for row, di, Di, si, scale in pending:
        row_update(state, cfg, dim, row, di, Di, scale)
    if cfg["noise_added"]:
        for row, di, Di, si, scale in pending:
            update_sigma_param(state, cfg, si, scale)
    return [p[0] for p in pending]


def update_hyper_parameter(state, cfg, dim):
    """update_hyper_parameter (:146-160): pSigma[dim] = sum_rows (tr S + m.m) / (n_d D)."""
    p = state["params"][dim]
    m = p[:, 0, :]
    if cfg["diag"]:
        tr = p[:, 1, :].sum()
    else:
        tr = np.trace(p[:, 1:, :], axis1=1, axis2=2).sum()
    state["pSigma"][dim] = (tr + np.sum(m * m)) / (p.shape[0] * cfg["D"])


def check_stop_cond(state):
    """check_stop_cond (:967-984)."""
    d_mean = max(float(a[:, 0].max()) for a in state["norm_changes"])
    d_cov = max(float(a[:, 1].max()) for a in state["norm_changes"])
    return max(d_mean, 0), max(d_cov, 0)


def predict_mean(state, entries):
    """predict_entry for the normal likelihood (:1039-1044): sum_D prod_modes m."""
    u = np.ones((entries.shape[0], state["params"][0].shape[2]))
    for d in range(entries.shape[1]):
        u = u * state["params"][d][entries[:, d], 0, :]
    return u.sum(axis=1)


# --------------------------------------------------------------------------------------
# Report-path estimators
# --------------------------------------------------------------------------------------
def log_pdf(lik, y, f, s):
    """normal.py:29-30, bernoulli.py:38-39: log of the floored pdf; poisson.py:59-68: scipy's
    poisson.logpmf(y, poisson_link(f)) = xlogy(y, A) - gammaln(y + 1) - A, nan_to_num."""
    if lik == LIK_POISSON:
        A = poisson_link(f)
        with np.errstate(all="ignore"):
            xl = np.where(y == 0, 0.0, y * np.log(A))
            return np.nan_to_num(xl - math.lgamma(float(y) + 1.0) - A)
    return np.log(pdf(lik, y, f, s))


def link_fun(lik, f):
    """SSVI_TF_interface.py:60-75."""
    if lik == LIK_NORMAL:
        return f
    if lik == LIK_BERNOULLI:
        return sigmoid(f)
    return poisson_link(f)


EST_EXPECTED_OBS, EST_LOG_PDF, EST_PDF, EST_ROBUST_VLB = 0, 1, 2, 3


def estimate_entries(state, cfg, what, entries, ys, pool, slot0=0):
    """The shared sampling pattern of estimate_expected_observation_via_sampling (:1168-1189),
    estimate_expectation_term_vlb (:889-908) and estimate_negative_log_likelihood (:939-964): for EVERY mode of
    the entry k draws x = m + L z, ms_k = sum_D prod_modes x; fs = ms (SSVI_TF_d) or
    normal(ms, rayleigh(w_sigma^2, (k,)), size=(k2, k)); mean over all draws of link(fs) / log_pdf(y, fs) /
    pdf(y, fs).  SSVI_TF_robust overrides the VLB term (SSVI_TF_robust.py:234-262): the link is applied BEFORE the
    pdf, the pdf is averaged over the first-layer draws for each second-layer draw, and the logs are averaged:
    mean_i log mean_k pdf(y, link(fs[i][k])) (EST_ROBUST_VLB).  pool: Tape of P slots -- entry e reads slot
    (slot0 + e) mod P, mode 0 z_own, mode d >= 1 z_other[:, d-1]."""
    P = pool.z_own.shape[0]
    N = entries.shape[1]
    lik, s = cfg["likelihood"], cfg["likelihood_param"]
    out = np.zeros(len(entries))
    for e, entry in enumerate(entries):
        sl = (slot0 + e) % P
        prod = np.ones((cfg["k1"], cfg["D"]))
        for d in range(N):
            m, S = get_vector_distribution(state, cfg, d, entry[d])
            z = pool.z_own[sl] if d == 0 else pool.z_other[sl, d - 1]
            prod = prod * (m + z @ _chol(cfg, S).T)
        ms = prod.sum(axis=1)
        if cfg["noise_added"]:
            ws = state["w_sigma"] ** 2 * pool.r_w[sl]
            fs = ms[None, :] + ws[None, :] * pool.z_f[sl]
        else:
            fs = ms
        y = None if ys is None else ys[e]
        if what == EST_EXPECTED_OBS:
            out[e] = np.mean(link_fun(lik, fs))
        elif what == EST_LOG_PDF:
            out[e] = np.mean(log_pdf(lik, y, fs, s))
        elif what == EST_PDF:
            out[e] = np.mean(pdf(lik, y, fs, s))
        else:
            out[e] = np.mean(np.log(np.mean(pdf(lik, y, link_fun(lik, fs), s), axis=1)))
    return out


def estimate_vlb(state, cfg, entries, ys, pool):
    """estimate_vlb (:874-883): (sum of the expectation terms, KL)."""
    what = EST_ROBUST_VLB if cfg["model"] == "robust" else EST_LOG_PDF
    return float(estimate_entries(state, cfg, what, entries, ys, pool).sum()), estimate_KL_divergence(state, cfg)


def estimate_negative_log_likelihood(state, cfg, entries, ys, pool):
    """(:939-964): -sum_e log mean pdf."""
    return float(-np.log(estimate_entries(state, cfg, EST_PDF, entries, ys, pool)).sum())


def estimate_KL_divergence(state, cfg):
    """(:911-936).  The loop OVERWRITES KL (:923-925), so the value returned is the divergence of the LAST row of
    the LAST mode from N(pmu, pSigma[dim] I) -- reproduced as is."""
    dim = len(state["params"]) - 1
    col = state["params"][dim].shape[0] - 1
    m, S = get_vector_distribution(state, cfg, dim, col)
    S0 = np.diag(S) if cfg["diag"] else S
    D = cfg["D"]
    sigma = state["pSigma"][dim]
    dm = state["pmu"] - m
    return float(0.5 * (np.trace(S0) / sigma + dm @ dm / sigma - D + np.log(sigma ** D / np.linalg.det(S0))))


# --------------------------------------------------------------------------------------
# Index (Tensor/Tensor.py:76-83, 152-169, 251-258): per-mode CSR in COO insertion order
# --------------------------------------------------------------------------------------
def build_index(coords, dims):
    """rowptr[d] (n_d+1,) int64 and perm[d] (nnz,) int64 such that
    perm[d][rowptr[d][i]:rowptr[d][i+1]] lists the entries touching row i of mode d in COO order
    (== the append order of observed_by_id[d][i])."""
    coords = np.asarray(coords)
    rowptr, perm = [], []
    for d, n in enumerate(dims):
        col = coords[:, d]
        perm.append(np.argsort(col, kind="stable").astype(np.int64))
        cnt = np.bincount(col, minlength=n)
        rp = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(cnt, out=rp[1:])
        rowptr.append(rp)
    return rowptr, perm


class Tape(object):
    """Travelling twin of oracle/ref_shim.py:NoiseTape (no reference import needed)."""

    def __init__(self, z_other, z_own, r_w=None, z_f=None, sub_idx=None):
        self.z_other, self.z_own, self.r_w, self.z_f, self.sub_idx = z_other, z_own, r_w, z_f, sub_idx


def make_tape(rng, cfg, N, B, n_obs=None):
    """Draw the standard variates one row minibatch consumes (B entries)."""
    k1, k2, D = cfg["k1"], cfg["k2"], cfg["D"]
    sub = None
    if n_obs is not None and n_obs > cfg["batch_size"]:
        sub = rng.choice(n_obs, cfg["batch_size"], replace=False)      # unsorted, like np.random.choice
        B = cfg["batch_size"]
    f32 = lambda a: a.astype(np.float32).astype(np.float64)   # noqa: E731  (float32-representable,
    # so the same tape feeds the device's f32 and f64 noise pools bit-identically)
    return Tape(f32(rng.standard_normal((B, N - 1, k1, D))), f32(rng.standard_normal((B, k1, D))),
                f32(rng.rayleigh(1.0, (B, k1))), f32(rng.standard_normal((B, k2, k1))), sub)
