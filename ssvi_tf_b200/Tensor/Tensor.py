"""Data containers (mirror of Tensor/Tensor.py): COO train/test/valid entries + the per-mode inverted
index.  The reference keeps the index as nested Python lists observed_by_id[dim][row] = [(coord, val)];
here it is a per-mode CSR (rowptr, perm) over the COO arrays, in the same (insertion) order, and the
list view is materialised lazily per row by find_observed_ui.
"""
import time

import numpy as np

from ..engine import build_csr


class _ObservedView(object):
    """observed_by_id[dim][row] -> list of (coord, val), built on demand from the CSR."""

    class _Mode(object):
        def __init__(self, tensor, dim):
            self.t, self.dim = tensor, dim

        def __len__(self):
            return self.t.dims[self.dim]

        def __getitem__(self, row):
            return self.t.find_observed_ui(self.dim, row)

    def __init__(self, tensor):
        self.t = tensor

    def __len__(self):
        return len(self.t.dims)

    def __getitem__(self, dim):
        return _ObservedView._Mode(self.t, dim)


class Tensor(object):
    def __init__(self, datatype="real"):
        assert datatype in ["real", "ordinal", "count", "binary"]
        self.datatype = datatype
        self._rowptr = self._perm = None

    # ---------------------------------------------------------------------------------------------
    # index
    # ---------------------------------------------------------------------------------------------
    def _set_split(self, entries, vals, train_num, valid_num=0):
        entries = np.asarray(entries, dtype=np.int64).reshape(-1, len(self.dims))
        vals = np.asarray(vals, dtype=np.float64)
        self.train_entries, self.train_vals = entries[:train_num], vals[:train_num]
        self.valid_entries = entries[train_num:train_num + valid_num]
        self.valid_vals = vals[train_num:train_num + valid_num]
        self.test_entries, self.test_vals = entries[train_num + valid_num:], vals[train_num + valid_num:]
        self._build_index()

    def _build_index(self):
        """Per-mode CSR in COO insertion order == observed_by_id's append order
        (Tensor/Tensor.py:76-83, 152-169)."""
        self._rowptr, self._perm = build_csr(self.train_entries, self.dims)
        self.observed_by_id = _ObservedView(self)

    def index(self):
        return self._rowptr, self._perm

    def find_observed_ui(self, dim, i):
        """All train entries touching row i of mode dim, as the reference's list of (coord, val)
        (Tensor/Tensor.py:251-258)."""
        ids = self._perm[dim][self._rowptr[dim][i]:self._rowptr[dim][i + 1]]
        return [(self.train_entries[e], self.train_vals[e]) for e in ids]

    def num_observed(self, dim, i):
        return int(self._rowptr[dim][i + 1] - self._rowptr[dim][i])

    # ---------------------------------------------------------------------------------------------
    # loaders / generators (host side, not on the timed path)
    # ---------------------------------------------------------------------------------------------
    def read_from_file(self, filename, train_size, validation_size, test_size):
        """CSV: header 'd1,d2,...' then 'i,j,k,val' per line; shuffled split (Tensor/Tensor.py:22-83)."""
        assert validation_size + train_size + test_size == 1
        with open(filename, "r") as f:
            header = f.readline().strip().split(",")
            self.dims = [int(x) for x in header]
            self.num_dimensions = len(header)
            data = np.loadtxt(f, delimiter=",", dtype=np.float64, ndmin=2)
        entries = data[:, :-1].astype(np.int64)
        vals = data[:, -1]
        if self.datatype == "count":
            self.min_count = min(getattr(self, "min_count", 100), int(vals.min()))
            self.max_count = max(getattr(self, "max_count", 0), int(vals.max()))
        total = len(vals)
        train_num = int(total * train_size)
        valid_num = int(total * validation_size)
        indices = np.arange(total)
        np.random.shuffle(indices)
        self._set_split(entries[indices], vals[indices], train_num, valid_num)

    def create_random_matrix(self, nrows, D, mean, cov):
        """Missing from the reference at this commit (called at real_tensor.py:19 etc., defined nowhere):
        nrows i.i.d. rows from N(mean, cov), shape (nrows, D) as compute_entry_value indexes it."""
        return np.random.multivariate_normal(mean, cov, size=nrows)

    def generate_hidden_matrices(self, means, covariances):
        raise NotImplementedError

    def data_link_fun(self, f):
        raise NotImplementedError

    def synthesize_data(self, dims, means, covariances, D=20, train=0.8, sparsity=1, noise=0.1, noise_ratio=True):
        """Synthetic CP tensor (Tensor/Tensor.py:88-125), vectorised."""
        print("Generating synthetic", self.datatype, "valued data ... ")
        start = time.time()
        self.dims = list(dims)
        self.D = D
        self.train = train
        self.noise = noise
        self.noise_ratio = noise_ratio
        self.matrices = self.generate_hidden_matrices(means, covariances)
        total = int(np.prod(dims))
        observed_num = int(total * sparsity)
        train_size = int(observed_num * train)
        self.observed_entries = self.generate_unique_coords(observed_num)
        self.observed_vals = self.compute_entry_values(self.observed_entries)
        self._set_split(self.observed_entries, self.observed_vals, train_size)
        print("Generating synthetic ", self.datatype, "valued data took: ", time.time() - start)
        if self.datatype == "count":
            print("max_count = ", self.max_count, " min count = ", self.min_count)

    def reduce_train_size(self, train_ratio):
        """Tensor/Tensor.py:132-150."""
        print("Using ", train_ratio, "of training data")
        n = int(self.train * train_ratio * len(self.observed_vals))
        self.train_entries = self.observed_entries[:n]
        self.train_vals = np.asarray(self.observed_vals[:n], dtype=np.float64)
        self._build_index()

    def compute_entry_values(self, entries):
        """compute_entry_value (Tensor/Tensor.py:171-189) for all entries at once."""
        entries = np.asarray(entries)
        u = np.ones((len(entries), self.D))
        for dim in range(len(self.dims)):
            u = u * self.matrices[dim][entries[:, dim], :]
        m = u.sum(axis=1)
        if self.noise != 0:
            stddev = self.noise * np.abs(m) if self.noise_ratio else np.full_like(m, self.noise)
            f = np.random.normal(m, stddev)
        else:
            f = m
        return self.data_link_fun(f)

    def compute_entry_value(self, entry):
        return self.compute_entry_values(np.asarray(entry)[None, :])[0]

    def generate_unique_coords(self, num):
        """num distinct cells, uniformly (Tensor/Tensor.py:260-281).  Exact choice-without-replacement
        for small index spaces; rejection sampling otherwise."""
        total = int(np.prod(self.dims))
        if total <= 50_000_000:
            idx = np.random.choice(total, num, replace=False)
        else:
            idx = np.unique(np.random.randint(0, total, size=int(num * 1.05) + 16, dtype=np.int64))
            while len(idx) < num:
                idx = np.unique(np.concatenate([idx, np.random.randint(0, total, size=num, dtype=np.int64)]))
            idx = np.random.permutation(idx)[:num]
        return np.stack(np.unravel_index(idx, self.dims), axis=1).astype(np.int64)

    def actual_value(self, m):
        if self.datatype == "real":
            return m
        if self.datatype == "binary":
            return 1 if m >= self.binary_cutoff else -1
        raise NotImplementedError
