"""Variational posterior store (mirror of Model/TF_Models.py), backed by device-resident tensors.

Layout is the reference's: per mode a dense array (n_rows, D+1, D) -- row 0 the mean, rows 1..D the
full covariance -- or (n_rows, 2, D) for a diagonal covariance (Model/TF_Models.py:53-93).  The
arrays live in HBM (torch CUDA tensors owned by the Engine); get/update_vector_distribution move one
row between host and device.
"""
import numpy as np
import torch


class Posterior(object):
    def __init__(self, dims, D, initMean=None, initCov=None, randstart=True):
        self.dims = dims
        self.D = D
        self.randstart = randstart
        if initMean is None or initCov is None:          # TF_Models.py:14-16
            initMean = np.ones((D,))
            initCov = np.eye(D)
        self.initMean = np.asarray(initMean, dtype=np.float64)
        self.initCov = np.asarray(initCov, dtype=np.float64)
        self.params = None        # list of device tensors, set by attach()
        self._on_write = None

    # -- host-side initial values ---------------------------------------------------------------
    def initial_values(self):
        """One numpy array per mode.  Random start draws each mean from N(initMean, initCov)
        (TF_Models.py:53-62, 77-86) with numpy's global RNG, one multivariate draw per mode."""
        out = []
        for n in self.dims:
            p = np.zeros(self._shape(n))
            if self.randstart:
                p[:, 0, :] = np.random.multivariate_normal(self.initMean, self.initCov, size=n)
            else:
                p[:, 0, :] = self.initMean
            self._fill_cov(p)
            out.append(p)
        return out

    def attach(self, device_params, on_write):
        """Bind to the Engine's tensors; on_write(dim, row) refreshes the sampler cache row."""
        self.params = device_params
        self._on_write = on_write

    def reset(self):
        for d, p in enumerate(self.initial_values()):
            self.params[d].copy_(torch.as_tensor(p))
            self._on_write(d, None)

    def save_mean_params(self, dim, filename):
        np.savetxt(filename, self.params[dim][:, 0, :].cpu().numpy(), delimiter=",")

    def save_mean_parameters(self, dim, mean_file):
        np.save(mean_file, self.params[dim][:, 0, :].cpu().numpy())

    def save_cov_parameters(self, dim, cov_file):
        np.save(cov_file, self.params[dim][:, 1:, :].cpu().numpy())


class Posterior_Full_Covariance(Posterior):
    def _shape(self, n):
        return (n, self.D + 1, self.D)

    def _fill_cov(self, p):
        p[:, 1:, :] = self.initCov

    def get_vector_distribution(self, dim, i):
        row = self.params[dim][i].cpu().numpy()
        return row[0, :], row[1:, :]

    def update_vector_distribution(self, dim, i, m_next, S_next):
        p = self.params[dim]
        p[i, 0, :] = torch.as_tensor(np.asarray(m_next, dtype=np.float64), device=p.device)
        p[i, 1:, :] = torch.as_tensor(np.asarray(S_next, dtype=np.float64), device=p.device)
        self._on_write(dim, i)


class Posterior_Diag_Covariance(Posterior):
    def _shape(self, n):
        return (n, 2, self.D)

    def _fill_cov(self, p):
        p[:, 1, :] = np.diag(self.initCov)

    def get_vector_distribution(self, dim, i):
        row = self.params[dim][i].cpu().numpy()
        return row[0, :], row[1, :]

    def update_vector_distribution(self, dim, i, m_next, S_next):
        p = self.params[dim]
        p[i, 0, :] = torch.as_tensor(np.asarray(m_next, dtype=np.float64), device=p.device)
        p[i, 1, :] = torch.as_tensor(np.asarray(S_next, dtype=np.float64), device=p.device)
        self._on_write(dim, i)
