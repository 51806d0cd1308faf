"""A NumPy stand-in for the handful of CasADi names the reference GP layer uses  --  TEST INFRASTRUCTURE ONLY.

casadi is not installed (and not installable) in the build image, so the reference's `covSE_single`,
`covSE_vectorized` and `gpytorch_predict2casadi` (gpmpc/gp.py:12-21, 72-85) could never be executed.  They are
plain arithmetic over a `cs` namespace: `cs.sum1`, `cs.exp`, `cs.repmat`, `cs.SX.sym`, `cs.Function`, `.T`, `@`
and the elementwise operators.  This module provides exactly those over numpy float64 arrays with CasADi's
conventions (a vector symbol is a (d,1) column; `sum1` sums over rows and keeps a (1,n) row; a binary operation
between a (d,1) column and a (d,n) matrix repeats the column, which numpy broadcasting reproduces), so that
  * tests/golden/make_golden_covse.py can run the UNMODIFIED reference functions and commit their outputs, and
  * the import-guarded CasADi twins in gpmpc_b200/gp.py execute at least once in the test-suite.

Symbols are lazy expressions (a closure over an environment) so `Function(name, [z], [expr], in, out)` can be
called repeatedly with numeric arguments, by position or by keyword, and returns a dict like casadi does.
"""

from __future__ import annotations

import numpy as np


class Expr:
    """Lazy numpy expression: `ev(env)` evaluates with env = {id(symbol): ndarray}."""

    __array_priority__ = 1000   # numpy operands defer to Expr.__r*__

    def __init__(self, ev, shape):
        self._ev, self.shape = ev, tuple(shape)

    @staticmethod
    def wrap(v):
        if isinstance(v, Expr):
            return v
        a = np.asarray(v, dtype=np.float64)
        return Expr(lambda env, a=a: a, a.shape)

    def _bin(self, other, op, swap=False):
        o = Expr.wrap(other)
        a, b = (o, self) if swap else (self, o)
        shape = np.broadcast_shapes(a.shape, b.shape)
        return Expr(lambda env: op(a._ev(env), b._ev(env)), shape)

    def __add__(self, o): return self._bin(o, np.add)
    def __radd__(self, o): return self._bin(o, np.add, True)
    def __sub__(self, o): return self._bin(o, np.subtract)
    def __rsub__(self, o): return self._bin(o, np.subtract, True)
    def __mul__(self, o): return self._bin(o, np.multiply)
    def __rmul__(self, o): return self._bin(o, np.multiply, True)
    def __truediv__(self, o): return self._bin(o, np.divide)
    def __rtruediv__(self, o): return self._bin(o, np.divide, True)
    def __pow__(self, o): return self._bin(o, np.power)
    def __neg__(self): return Expr(lambda env: -self._ev(env), self.shape)

    def __matmul__(self, o):
        o = Expr.wrap(o)
        shape = (self.shape[0],) + tuple(o.shape[1:])
        return Expr(lambda env: self._ev(env) @ o._ev(env), shape)

    def __rmatmul__(self, o):
        return Expr.wrap(o).__matmul__(self)

    @property
    def T(self):
        return Expr(lambda env: self._ev(env).T, self.shape[::-1])


class SX(Expr):
    @staticmethod
    def sym(name, n=1, m=1):
        s = Expr(None, (n, m))
        s.name = name
        s._ev = lambda env, key=id(s): env[key]
        return s


MX = SX


def _lift(fn, shape_fn):
    def f(x, *args):
        if isinstance(x, Expr):
            return Expr(lambda env: fn(x._ev(env), *args), shape_fn(x.shape, *args))
        return fn(np.asarray(x, dtype=np.float64), *args)
    return f


def _sum1(a):
    a = np.atleast_2d(a)
    return a.sum(axis=0, keepdims=True)


def _repmat(a, n, m=1):
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a[:, None]          # casadi vectors are columns
    return np.tile(a, (n, m))


sum1 = _lift(_sum1, lambda s: (1, s[1] if len(s) > 1 else 1))
exp = _lift(np.exp, lambda s: s)
repmat = _lift(_repmat, lambda s, n, m=1: (s[0] * n, (s[1] if len(s) > 1 else 1) * m))


class Function:
    """cs.Function(name, [inputs], [outputs], in_names, out_names); call by position or keyword -> dict / value."""

    def __init__(self, name, ins, outs, in_names=None, out_names=None):
        self.name_, self.ins, self.outs = name, list(ins), [Expr.wrap(o) for o in outs]
        self.in_names = list(in_names) if in_names else [f"i{k}" for k in range(len(ins))]
        self.out_names = list(out_names) if out_names else [f"o{k}" for k in range(len(outs))]

    def name(self):
        return self.name_

    def __call__(self, *args, **kw):
        vals = list(args) + [kw[n] for n in self.in_names[len(args):]]
        assert len(vals) == len(self.ins), f"{self.name_}: expected {len(self.ins)} inputs"
        if any(isinstance(v, Expr) for v in vals):
            # symbolic call (K_xz(z=z) at gpmpc/gp.py:85): substitute the arguments for the inputs, stay lazy
            subs = [(id(sym), Expr.wrap(v)) for sym, v in zip(self.ins, vals)]

            def substituted(o):
                return Expr(lambda env: o._ev({**env, **{k: e._ev(env).reshape(-1, 1) if e._ev(env).ndim == 1
                                                         else e._ev(env) for k, e in subs}}), o.shape)

            res = {n: substituted(o) for n, o in zip(self.out_names, self.outs)}
        else:
            env = {}
            for sym, v in zip(self.ins, vals):
                a = np.asarray(v, dtype=np.float64)
                env[id(sym)] = a.reshape(sym.shape)
            res = {n: np.asarray(o._ev(env)) for n, o in zip(self.out_names, self.outs)}
        if kw:
            return res
        out = [res[n] for n in self.out_names]
        return out[0] if len(out) == 1 else out


class Sparsity:
    """Only the dense pattern is needed by the Callback bridge: a (rows, cols) shape."""

    def __init__(self, n, m):
        self.shape = (n, m)

    @staticmethod
    def dense(n, m=1):
        return Sparsity(n, m)


class Callback:
    """casadi.Callback's protocol as far as gpmpc_b200.gp.gp_casadi_callback uses it: construct(name, opts) finalises,
    calling the object evaluates `eval` on numeric inputs shaped by get_sparsity_in, `jacobian()` returns the Function
    get_jacobian builds (CasADi calls it with the nominal inputs followed by the nominal outputs)."""

    def __init__(self):
        self._name = None

    def construct(self, name, opts=None):
        self._name, self._opts = name, dict(opts or {})

    def name(self):
        return self._name

    def __call__(self, *args, **kw):
        n_in = self.get_n_in()
        if kw:
            names = [self.get_name_in(i) if hasattr(self, "get_name_in") else f"i{i}" for i in range(n_in)]
            args = [kw[n] for n in names]
        vals = [np.asarray(a, dtype=np.float64).reshape(self.get_sparsity_in(i).shape) for i, a in enumerate(args)]
        out = [np.asarray(o, dtype=np.float64).reshape(self.get_sparsity_out(i).shape) for i, o in enumerate(self.eval(vals))]
        if kw:
            onames = [self.get_name_out(i) if hasattr(self, "get_name_out") else f"o{i}" for i in range(len(out))]
            return dict(zip(onames, out))
        return out[0] if len(out) == 1 else out

    def jacobian(self):
        assert self.has_jacobian()
        return self.get_jacobian("jac_" + self._name, None, None, {})
