"""Ordinary polynomial regression with the regression matrix built on the device.

Mirror of ``src/minterpy/extras/regression/ordinary_regression.py`` for the two bases that
sit on the hot path:

* Newton basis: the regression matrix IS the (P x N) Newton-monomial matrix
  (``compute_regression_matrix`` :259-262 -> ``eval_newton_monomials``) -- the
  write-bandwidth-bound kernel ``mpx_monomials``;
* Lagrange basis (the reference's default): the N Lagrange monomials are Newton polynomials
  whose coefficients are the columns of the Lagrange->Newton matrix, so the regression matrix
  is ``eval_newton_polynomials(xx, L2N, ...)`` with k = N coefficient columns (:241-257) --
  the tensor-core kernel behind ``mpx_eval`` with the widest k the library sees.  The L2N
  matrix itself is the dds kernel applied to the identity.

The least-squares solve and the error measures are dense LAPACK-type steps in the reference
(SciPy/NumPy) and stay library calls here: SciPy/NumPy for NumPy inputs, ``torch.linalg`` for
CUDA tensors (then nothing leaves the device).
"""
import numpy as np

from . import api
from .plan import _is_torch
from .polynomials import (Grid, LagrangePolynomial, LagrangeToNewton, MultiIndexSet,
                          NewtonPolynomial)

__all__ = ["OrdinaryRegression", "compute_regression_matrix"]


def compute_regression_matrix(basis_poly, xx):
    """(P x N) values of the basis monomials at ``xx`` (``ordinary_regression.py:221-275``)."""
    e = basis_poly.multi_index.exponents
    gp = basis_poly.grid.generating_points
    if isinstance(basis_poly, LagrangePolynomial):
        op = LagrangeToNewton(basis_poly).transformation_operator
        if _is_torch(xx):
            # the N x N Lagrange->Newton matrix never visits the host: identity and dds on
            # the device, then evaluation with k = N coefficient columns
            l2n = op.full_on_device(xx.device)
        else:
            l2n = op.array_repr_full
        rr = api.eval_newton_polynomials(xx, l2n, e, gp)
        # the reference's two edge cases: a single point, a single basis function
        if xx.shape[0] == 1:
            rr = rr[None, :]
        if rr.ndim == 1:
            rr = rr[:, None]
        return rr
    if isinstance(basis_poly, NewtonPolynomial):
        return api.eval_newton_monomials(xx, e, gp)
    raise TypeError(f"Polynomial {type(basis_poly)} is not supported!")


def _lin(xp_is_torch):
    if xp_is_torch:
        import torch
        return torch.linalg
    return np.linalg


def _diag_weights(weights):
    if weights is None:
        return None
    if weights.ndim == 1:
        if _is_torch(weights):
            import torch
            return torch.diag(weights)
        return np.diag(weights)
    return weights


def solve_least_squares(rr, yy, weights=None, lstsq_solver="lstsq", **kwargs):
    """``ordinary_regression.py:278-353``; solvers ``lstsq``, ``inv``, ``pinv``, ``qr``, or a
    callable ``solver(rr, yy, weights, **kwargs)``."""
    ww = _diag_weights(weights)
    if callable(lstsq_solver):
        return lstsq_solver(rr, yy, ww, **kwargs)
    name = lstsq_solver.lower()
    on_device = _is_torch(rr)
    la = _lin(on_device)
    if ww is not None:
        a, b = rr.T @ ww @ rr, rr.T @ ww @ yy
    else:
        a, b = rr, yy
    if name == "lstsq":
        if on_device:
            b2 = b[:, None] if b.ndim == 1 else b
            sol = la.lstsq(a, b2).solution
            return sol[:, 0] if b.ndim == 1 else sol
        import scipy.linalg
        return scipy.linalg.lstsq(a, b, **kwargs)[0]
    if name == "inv":
        if ww is None:
            return la.inv(rr.T @ rr) @ rr.T @ yy
        return la.inv(a) @ b
    if name == "pinv":
        return la.pinv(a, **kwargs) @ b
    if name == "qr":
        if ww is None and rr.shape[0] < rr.shape[1]:
            qq, r_ = la.qr(rr.T)
            return qq @ la.inv(r_.T) @ yy
        qq, r_ = la.qr(a)
        return la.solve(r_, qq.T @ b)
    raise NotImplementedError(f"{lstsq_solver} solver not implemented!")


def _to_float(v):
    return float(v.item()) if hasattr(v, "item") else float(v)


def compute_regfit_l1_error(rr, coeffs, yy):
    """Largest absolute residual, absolute and relative to std(yy) (:404-434)."""
    err = abs(yy - rr @ coeffs).max()
    std = yy.std(unbiased=False) if _is_torch(yy) else np.std(yy)
    return _to_float(err), _to_float(err / std)


def compute_regfit_l2_error(rr, coeffs, yy):
    """Mean squared residual, absolute and relative to var(yy) (:437-466)."""
    err = ((yy - rr @ coeffs) ** 2).mean()
    var = yy.var(unbiased=False) if _is_torch(yy) else np.var(yy)
    return _to_float(err), _to_float(err / var)


def compute_loocv_error(rr, coeffs, yy, weights=None):
    """Leave-one-out error from the hat-matrix diagonal (:356-401)."""
    on_device = _is_torch(rr)
    la = _lin(on_device)
    var = yy.var(unbiased=False) if on_device else np.var(yy)
    if rr.shape[0] < rr.shape[1]:
        return float("inf"), float("inf")
    ww = _diag_weights(weights)
    if ww is None:
        gram_inv = la.pinv(rr.T @ rr)
        hi = ((rr @ gram_inv) * rr).sum(1)          # diag(R G^+ R^T) without the P x P matrix
    else:
        gram_inv = la.pinv(rr.T @ ww @ rr)
        hi = ((rr @ gram_inv) * (ww.T @ rr)).sum(1)
    loo = ((yy - rr @ coeffs) / (1 - hi)) ** 2
    if on_device:
        import torch
        loo = torch.where(torch.isnan(loo), torch.full_like(loo, float("inf")), loo)
    else:
        loo[np.isnan(loo)] = np.inf
    err = loo.mean()
    return _to_float(err), _to_float(err / var)


class OrdinaryRegression:
    """``OrdinaryRegression(multi_index, grid, origin_poly)`` with ``fit`` / ``predict``
    (``ordinary_regression.py:26-218``)."""

    def __init__(self, multi_index=None, grid=None, origin_poly=LagrangePolynomial):
        if multi_index is None and grid is None:
            raise ValueError("Either multi-index set or grid must be specified!")
        if grid is None:
            if not isinstance(multi_index, MultiIndexSet):
                raise TypeError(f"Unexpected type {type(multi_index)} of multi-index set.")
            grid = Grid(multi_index)
        elif not isinstance(grid, Grid):
            raise TypeError(f"Unexpected type {type(grid)} of grid.")
        if multi_index is None:
            multi_index = grid.multi_index
        elif not isinstance(multi_index, MultiIndexSet):
            raise TypeError(f"Unexpected type {type(multi_index)} of multi-index set.")
        self._grid = grid
        self._multi_index = multi_index
        self._origin_poly = origin_poly(multi_index, None, grid=grid)
        self._regression_matrix = None
        self._coeffs = None
        self._loocv_error = None
        self._regfit_l1_error = None
        self._regfit_l2_error = None
        self._eval_poly = None

    multi_index = property(lambda self: self._multi_index)
    grid = property(lambda self: self._grid)
    origin_poly = property(lambda self: self._origin_poly)
    eval_poly = property(lambda self: self._eval_poly)
    coeffs = property(lambda self: self._coeffs)
    regression_matrix = property(lambda self: self._regression_matrix)
    loocv_error = property(lambda self: self._loocv_error)
    regfit_l1_error = property(lambda self: self._regfit_l1_error)
    regfit_l2_error = property(lambda self: self._regfit_l2_error)

    def get_regression_matrix(self, xx):
        return compute_regression_matrix(self._origin_poly, xx)

    def fit(self, xx, yy, weights=None, lstsq_solver="lstsq", **kwargs):
        rr = self.get_regression_matrix(xx)
        self._regression_matrix = rr
        self._coeffs = solve_least_squares(rr, yy, weights, lstsq_solver, **kwargs)
        self._regfit_l1_error = compute_regfit_l1_error(rr, self._coeffs, yy)
        self._regfit_l2_error = compute_regfit_l2_error(rr, self._coeffs, yy)
        self._loocv_error = compute_loocv_error(rr, self._coeffs, yy, weights)
        if isinstance(self._origin_poly, LagrangePolynomial):
            lag = LagrangePolynomial(self._multi_index, self._coeffs, grid=self._grid)
            self._eval_poly = LagrangeToNewton(lag)()
        else:
            self._eval_poly = NewtonPolynomial(self._multi_index, self._coeffs, grid=self._grid)

    def predict(self, xx):
        if self._eval_poly is None:
            raise TypeError("Ordinary regression model is not fitted, "
                            "regression polynomial can't be evaluated!")
        return self._eval_poly(xx)

    __call__ = predict
