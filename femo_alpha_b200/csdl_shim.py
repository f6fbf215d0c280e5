"""Minimal stand-in for the parts of `csdl_alpha` that femo_alpha's operator surface touches.

Used ONLY when `import csdl_alpha` fails (it is not installable here).  It gives the
operations in `csdl_alpha_opt/` the same base classes and calling protocol the reference
relies on (femo_alpha/csdl_alpha_opt/state_operation.py:8,47-131,134-220;
output_operation.py:6-69): `declare_input`, `create_output`,
`declare_derivative_parameters`, inline evaluation, and a reverse sweep that drives
`compute_derivatives` / `apply_inverse_jacobian` / `compute_jacvec_product` exactly the way
CSDL's simulator does.  It is an eager tape, not a graph compiler.
"""
from __future__ import annotations

import types

import numpy as np

_active_recorder = None


class Recorder:
    def __init__(self, inline=True, **kw):
        self.inline = inline
        self.nodes = []

    def start(self):
        global _active_recorder
        _active_recorder = self

    def stop(self):
        global _active_recorder
        if _active_recorder is self:
            _active_recorder = None


def _record(node):
    if _active_recorder is not None:
        _active_recorder.nodes.append(node)


class _Node:
    """One recorded operation: `forward()` recomputes outputs in place, `vjp(out_bars)` -> input bars."""

    def __init__(self, inputs, outputs, forward, vjp):
        self.inputs, self.outputs, self.forward, self.vjp = inputs, outputs, forward, vjp
        for o in outputs:
            o._node = self
        _record(self)


class Variable:
    def __init__(self, value=None, shape=None, name=None):
        if value is None:
            value = np.zeros(shape if shape is not None else (1,))
        v = np.array(value, dtype=np.float64)
        if shape is not None and v.shape != tuple(shape):
            v = np.broadcast_to(v, shape).copy()
        if v.ndim == 0:
            v = v.reshape(1)
        self.value = v
        self.names = [name] if name else []
        self._node = None
        self.is_design_variable = self.is_constraint = self.is_objective = False
        self.meta = {}

    @property
    def shape(self):
        return self.value.shape

    @property
    def name(self):
        return self.names[-1] if self.names else None

    def add_name(self, name):
        self.names.append(name)
        return self

    def set_as_design_variable(self, **kw):
        self.is_design_variable = True; self.meta.update(kw)

    def set_as_constraint(self, **kw):
        self.is_constraint = True; self.meta.update(kw)

    def set_as_objective(self, **kw):
        self.is_objective = True; self.meta.update(kw)

    # ---- recorded array ops ---------------------------------------------------
    def __getitem__(self, idx):
        if isinstance(idx, list):
            idx = np.asarray(idx)
        out = Variable(value=self.value[idx])

        def fwd():
            out.value = np.array(self.value[idx], dtype=np.float64)

        def vjp(bars):
            g = np.zeros_like(self.value)
            np.add.at(g, idx, bars[0])
            return [g]
        _Node([self], [out], fwd, vjp)
        return out

    def reshape(self, shape):
        return reshape(self, shape)

    def _binary(self, other, f, dfa, dfb):
        if not isinstance(other, Variable):
            c = np.asarray(other, dtype=np.float64)
            out = Variable(value=f(self.value, c))

            def fwd():
                out.value = np.asarray(f(self.value, c), dtype=np.float64)

            def vjp(bars):
                return [_unbroadcast(bars[0] * dfa(self.value, c), self.value.shape)]
            _Node([self], [out], fwd, vjp)
            return out
        out = Variable(value=f(self.value, other.value))

        def fwd():
            out.value = np.asarray(f(self.value, other.value), dtype=np.float64)

        def vjp(bars):
            return [_unbroadcast(bars[0] * dfa(self.value, other.value), self.value.shape),
                    _unbroadcast(bars[0] * dfb(self.value, other.value), other.value.shape)]
        _Node([self, other], [out], fwd, vjp)
        return out

    def __add__(self, o): return self._binary(o, lambda a, b: a + b, lambda a, b: 1.0, lambda a, b: 1.0)
    __radd__ = __add__
    def __sub__(self, o): return self._binary(o, lambda a, b: a - b, lambda a, b: 1.0, lambda a, b: -1.0)
    def __rsub__(self, o): return (-self) + o
    def __mul__(self, o): return self._binary(o, lambda a, b: a * b, lambda a, b: b, lambda a, b: a)
    __rmul__ = __mul__
    def __truediv__(self, o): return self._binary(o, lambda a, b: a / b, lambda a, b: 1.0 / b, lambda a, b: -a / b ** 2)
    def __rtruediv__(self, o): return self._binary(o, lambda a, b: b / a, lambda a, b: -b / a ** 2, None)
    def __neg__(self): return self * -1.0
    def __pow__(self, p): return self._binary(p, lambda a, b: a ** b, lambda a, b: b * a ** (b - 1), lambda a, b: a ** b * np.log(a))


def _unbroadcast(g, shape):
    g = np.asarray(g, dtype=np.float64)
    while g.ndim > len(shape):
        g = g.sum(axis=0)
    for i, s in enumerate(shape):
        if s == 1 and g.shape[i] != 1:
            g = g.sum(axis=i, keepdims=True)
    return g.reshape(shape)


class VariableGroup:
    """Attribute bag (csdl.VariableGroup); missing attributes read as None like the reference expects."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return None


def reshape(x, shape):
    shape = tuple(shape) if not isinstance(shape, int) else (shape,)
    out = Variable(value=x.value.reshape(shape))

    def fwd():
        out.value = x.value.reshape(shape).copy()
    _Node([x], [out], fwd, lambda bars: [bars[0].reshape(x.value.shape)])
    return out


def transpose(x):
    out = Variable(value=x.value.T.copy())

    def fwd():
        out.value = x.value.T.copy()
    _Node([x], [out], fwd, lambda bars: [bars[0].T])
    return out


def solve_linear(A, b):
    A = np.asarray(A)
    out = Variable(value=np.linalg.solve(A, b.value))

    def fwd():
        out.value = np.linalg.solve(A, b.value)
    _Node([b], [out], fwd, lambda bars: [np.linalg.solve(A.T, bars[0])])
    return out


def _matvec(M, x):
    out = Variable(value=np.asarray(M @ x.value).ravel())

    def fwd():
        out.value = np.asarray(M @ x.value).ravel()
    _Node([x], [out], fwd, lambda bars: [np.asarray(M.T @ bars[0]).reshape(x.value.shape)])
    return out


sparse = types.SimpleNamespace(matvec=_matvec)


def check_parameter(value, name, types=None, **kw):
    if types is not None and not isinstance(value, types):
        raise TypeError("parameter '%s' must be of type %s, got %s" % (name, types, type(value)))


# ---- custom operations --------------------------------------------------------------
class _CustomOperation:
    def __init__(self, *a, **kw):
        self._inputs, self._outputs = {}, {}

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        user_eval = cls.__dict__.get("evaluate")
        if user_eval is None:
            return

        def wrapped(self, *a, **k):
            self._inputs, self._outputs = {}, {}
            ret = user_eval(self, *a, **k)
            self._finalize()
            return ret
        wrapped.__doc__ = user_eval.__doc__
        cls.evaluate = wrapped

    def declare_input(self, name, var):
        if not isinstance(var, Variable):
            var = Variable(value=var)
        self._inputs[name] = var

    def create_output(self, name, shape):
        shape = tuple(shape) if not isinstance(shape, int) else (shape,)
        v = Variable(value=np.zeros(shape))
        self._outputs[name] = v
        return v

    def declare_derivative_parameters(self, of, wrt, dependent=True, **kw):
        pass

    def _vals(self):
        return {k: v.value for k, v in self._inputs.items()}


class CustomExplicitOperation(_CustomOperation):
    def _finalize(self):
        ins, outs = dict(self._inputs), dict(self._outputs)

        def fwd():
            ov = {}
            self.compute({k: v.value for k, v in ins.items()}, ov)
            for k, v in outs.items():
                v.value = np.asarray(ov[k], dtype=np.float64).reshape(v.value.shape)

        def vjp(bars):
            derivs = {}
            self.compute_derivatives({k: v.value for k, v in ins.items()}, {k: v.value for k, v in outs.items()}, derivs)
            res = []
            for kin, vin in ins.items():
                g = np.zeros(vin.value.size)
                for (kout, vout), bar in zip(outs.items(), bars):
                    if (kout, kin) in derivs and bar is not None:
                        J = np.asarray(derivs[kout, kin], dtype=np.float64).reshape(vout.value.size, vin.value.size)
                        g += J.T @ np.asarray(bar).ravel()
                res.append(g.reshape(vin.value.shape))
            return res
        fwd()
        _Node(list(ins.values()), list(outs.values()), fwd, vjp)


class CustomImplicitOperation(_CustomOperation):
    def _finalize(self):
        ins, outs = dict(self._inputs), dict(self._outputs)

        def fwd():
            ov = {k: v.value for k, v in outs.items()}
            self.solve_residual_equations({k: v.value for k, v in ins.items()}, ov)
            for k, v in outs.items():
                v.value = np.asarray(ov[k], dtype=np.float64).reshape(v.value.shape)

        def vjp(bars):
            iv = {k: v.value for k, v in ins.items()}
            ov = {k: v.value for k, v in outs.items()}
            d_out = {k: (np.zeros(v.value.shape) if b is None else np.asarray(b, dtype=np.float64).reshape(v.value.shape))
                     for (k, v), b in zip(outs.items(), bars)}
            d_res = {}
            self.apply_inverse_jacobian(iv, ov, d_out, d_res, "rev")
            d_in = {k: np.zeros(v.value.shape) for k, v in ins.items()}
            self.compute_jacvec_product(iv, ov, d_in, d_out, d_res, "rev")
            return [-d_in[k] for k in ins]                     # dy/dx = -(dR/dy)^-1 dR/dx
        fwd()
        _Node(list(ins.values()), list(outs.values()), fwd, vjp)


# ---- simulator: replay + reverse sweep (csdl.experimental.PySimulator subset) ----------
class PySimulator:
    def __init__(self, recorder):
        self.recorder = recorder

    def run(self):
        for n in self.recorder.nodes:
            n.forward()

    def compute_totals(self, ofs, wrts):
        """d of / d wrt for every pair, by one reverse sweep per scalar entry of each `of`."""
        if isinstance(ofs, Variable): ofs = [ofs]
        if isinstance(wrts, Variable): wrts = [wrts]
        totals = {}
        for of in ofs:
            rows = {id(w): np.zeros((of.value.size, w.value.size)) for w in wrts}
            for i in range(of.value.size):
                seed = np.zeros(of.value.size); seed[i] = 1.0
                bars = self._sweep(of, seed.reshape(of.value.shape))
                for w in wrts:
                    if id(w) in bars:
                        rows[id(w)][i] = bars[id(w)].ravel()
            for w in wrts:
                totals[of, w] = rows[id(w)]
        return totals

    def _sweep(self, of, seed):
        bars = {id(of): seed}
        for n in reversed(self.recorder.nodes):
            ob = [bars.get(id(o)) for o in n.outputs]
            if all(b is None for b in ob):
                continue
            ob = [b if b is not None else None for b in ob]
            ib = n.vjp([np.zeros(o.value.shape) if b is None else b for o, b in zip(n.outputs, ob)])
            for v, g in zip(n.inputs, ib):
                if g is None:
                    continue
                bars[id(v)] = g if id(v) not in bars else bars[id(v)] + g
        return bars

    def check_totals(self, ofs=None, wrts=None, step_size=1e-6, raise_on_error=False, verbose=True):
        """Adjoint totals against central finite differences (reference usage:
        examples/advanced_examples/simple_shell_opt/ex_simple_shell_opt.py:109-112)."""
        if isinstance(ofs, Variable): ofs = [ofs]
        if isinstance(wrts, Variable): wrts = [wrts]
        analytic = self.compute_totals(ofs, wrts)
        report = {}
        for w in wrts:
            base = w.value.copy()
            fd = {id(of): np.zeros((of.value.size, base.size)) for of in ofs}
            for j in range(base.size):
                for sgn in (+1, -1):
                    w.value = base.copy(); w.value.flat[j] += sgn * step_size
                    self.run()
                    for of in ofs:
                        fd[id(of)][:, j] += sgn * of.value.ravel() / (2 * step_size)
            w.value = base
            self.run()
            for of in ofs:
                a, f = analytic[of, w], fd[id(of)]
                err = np.linalg.norm(a - f) / max(np.linalg.norm(f), 1e-300)
                report[of, w] = dict(analytical=a, fd=f, relative_error=err)
                if verbose:
                    print("check_totals d(%s)/d(%s): |analytic|=%.6e |fd|=%.6e rel.err=%.3e"
                          % (of.name, w.name, np.linalg.norm(a), np.linalg.norm(f), err))
        return report


experimental = types.SimpleNamespace(CustomImplicitOperation=CustomImplicitOperation, PySimulator=PySimulator)
