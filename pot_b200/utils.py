"""Result containers with the reference's object shapes.

OTResult    mirrors ot.utils.OTResult  (ot/utils.py:1127-1422): same property names
            (`plan`, `potentials`, `value`, `value_linear`, `marginals`, `lazy_plan`, `log`, ...).
LazyTensor  mirrors ot.utils.LazyTensor (ot/utils.py:1425-1492): an (n, m) tensor defined by a function of
            index/slice arguments; named kwargs become attributes.

Both are plain host-side containers; the arrays they hold are torch CUDA tensors produced by libb2ot.so.
"""

import torch

__all__ = ["OTResult", "LazyTensor"]


class LazyTensor:
    """A tensor that is never stored: indexing evaluates `getitem(*index, **kwargs)` for the requested block.

    Behaviour follows ot.utils.LazyTensor (ot/utils.py:1425-1492): `shape` / `ndim`, every keyword argument is also
    an attribute, an index shorter than `ndim` is completed with full slices, anything but int / slice / tuple
    indices raises NotImplementedError, and the repr lists the attribute names.
    """

    def __init__(self, shape, getitem, **kwargs):
        self.shape, self.ndim = shape, len(shape)
        self._getitem, self.kwargs = getitem, kwargs
        self.__dict__.update(kwargs)

    def __getitem__(self, key):
        if isinstance(key, (int, slice)):
            key = (key,)
        elif not isinstance(key, tuple):
            raise NotImplementedError("Only integer, slice, and tuple indexing is supported")
        index = key + (slice(None),) * (self.ndim - len(key))
        return self._getitem(*index, **self.kwargs)

    def __repr__(self):
        return f"LazyTensor(shape={self.shape},attributes=({','.join(self.kwargs)}))"


class OTResult:
    """Container of an OT solve, property-compatible with the reference's OTResult."""

    def __init__(self, potentials=None, value=None, value_linear=None, value_quad=None, plan=None, log=None,
                 backend=None, sparse_plan=None, lazy_plan=None, status=None, batch_size=100):
        self._potentials = potentials
        self._value = value
        self._value_linear = value_linear
        self._value_quad = value_quad
        self._plan = plan
        self._log = log
        self._sparse_plan = sparse_plan
        self._lazy_plan = lazy_plan
        self._backend = backend if backend is not None else "torch-cuda (libb2ot)"
        self._status = status
        self._batch_size = batch_size

    def __repr__(self):
        s = "OTResult("
        if self._value is not None:
            s += "value={},".format(self._value)
        if self._value_linear is not None:
            s += "value_linear={},".format(self._value_linear)
        if self._plan is not None:
            s += "plan={}(shape={}),".format(self._plan.__class__.__name__, tuple(self._plan.shape))
        if self._lazy_plan is not None:
            s += "lazy_plan={}(shape={}),".format(self._lazy_plan.__class__.__name__, self._lazy_plan.shape)
        return (s[:-1] if s[-1] != "(" else s) + ")"

    @property
    def potentials(self):
        return self._potentials

    @property
    def potential_a(self):
        return self._potentials[0] if self._potentials is not None else None

    @property
    def potential_b(self):
        return self._potentials[1] if self._potentials is not None else None

    @property
    def plan(self):
        return self._plan

    @property
    def sparse_plan(self):
        if self._sparse_plan is not None:
            return self._sparse_plan
        return self._plan.to_sparse_csr() if self._plan is not None else None

    @property
    def lazy_plan(self):
        return self._lazy_plan

    @property
    def value(self):
        return self._value

    @property
    def value_linear(self):
        return self._value_linear

    @property
    def value_quad(self):
        return self._value_quad

    @property
    def marginals(self):
        if self._plan is not None or self._lazy_plan is not None:
            return self.marginal_a, self.marginal_b
        return None

    @property
    def marginal_a(self):
        if self._plan is not None:
            return torch.sum(self._plan, -1)
        if self._lazy_plan is not None:
            lp, bs = self._lazy_plan, self._batch_size
            return torch.cat([torch.sum(lp[i:i + bs, :], 1) for i in range(0, lp.shape[0], bs)])
        return None

    @property
    def marginal_b(self):
        if self._plan is not None:
            return torch.sum(self._plan, -2)
        if self._lazy_plan is not None:
            lp, bs = self._lazy_plan, self._batch_size
            return torch.cat([torch.sum(lp[:, j:j + bs], 0) for j in range(0, lp.shape[1], bs)])
        return None

    @property
    def status(self):
        return self._status

    @property
    def log(self):
        return self._log

    @property
    def a_to_b(self):
        raise NotImplementedError()

    @property
    def b_to_a(self):
        raise NotImplementedError()

    @property
    def citation(self):
        return "POT library: Flamary et al., POT: Python Optimal Transport, JMLR 22(78):1-8, 2021."
