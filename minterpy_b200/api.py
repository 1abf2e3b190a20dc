"""Python mirror of the reference's hot-path callables, backed by the CUDA library.

Same names, argument order, shape rules and exceptions as the reference snapshot:

* ``eval_newton_polynomials``  -- ``src/minterpy/utils.py:247-343``
* ``eval_newton_monomials``    -- ``src/minterpy/utils.py:177-244``
* ``dds``                      -- ``src/minterpy/dds.py:794-826``

Inputs may be NumPy arrays (results are NumPy arrays; the library moves the data) or CUDA
``torch`` tensors (zero copy, results are CUDA tensors).  The input rules below restate
``src/minterpy/core/verification.py:48-163``.
"""
import numpy as np

from .plan import _is_torch, get_plan

# The reference runs with DEBUG = True (src/minterpy/global_settings.py:12), which turns the
# dtype checks on for every evaluation; the mirror keeps that default.
DEBUG = True


def _dtype_name(a):
    return str(a.dtype).replace("torch.", "")


def _require_dtype(name, a, expected):
    # verification.py:127-137
    if _dtype_name(a) != expected:
        raise TypeError(f"expected dtype {expected} for {name} but the dtype is {_dtype_name(a)}")


def rectify_query_points(x, m):
    """(nr_points, x as (nr_points, m)) -- verification.py:48-85."""
    if x.ndim == 1:
        if m == 1:
            nr_points = x.shape[0]
        else:
            if x.shape[0] != m:
                raise ValueError(
                    f"points x given as vector of shape {tuple(x.shape)} (1D). "
                    f"detected dimensionality of the exponents however is {m}")
            nr_points = 1
        x = x.reshape(nr_points, m)
    else:
        nr_points, m_points = x.shape
        if m != m_points:
            raise ValueError(f"the dimensionality of the input points {m_points} "
                             f"does not match the polynomial dimensionality {m}")
    return nr_points, x


def rectify_eval_input(x, coefficients, exponents, verify_input):
    """verification.py:88-139."""
    N, m = exponents.shape
    if N == 0:
        raise ValueError("at least 1 monomial must be given")
    nr_points, x = rectify_query_points(x, m)
    if coefficients.ndim == 1:
        nr_polynomials = 1
        coefficients = coefficients.reshape(N, 1)
    else:
        N_coeffs, nr_polynomials = coefficients.shape
        if N != N_coeffs:
            raise ValueError(
                f"the coefficient amount {N_coeffs} does not match the amount of monomials {N}")
    if verify_input:
        _require_dtype("x", x, "float64")
        _require_dtype("coefficients", coefficients, "float64")
        _require_dtype("exponents", exponents, "int64")
    return N, coefficients, m, nr_points, nr_polynomials, x


def convert_eval_output(results):
    """Squeeze, but never down to 0-D -- verification.py:142-163."""
    out = results.squeeze()
    if out.ndim == 0:
        out = out.reshape(1)
    return out


def _as_host_int(exponents):
    if _is_torch(exponents):
        exponents = exponents.cpu().numpy()
    return np.asarray(exponents)


def _as_host_float(a):
    if _is_torch(a):
        a = a.cpu().numpy()
    return np.asarray(a)


def eval_newton_polynomials(xx, coefficients, exponents, generating_points,
                            verify_input=False, batch_size=None, devices=None):
    """Evaluate polynomial(s) in Newton form at the query points ``xx``.

    ``batch_size`` is accepted for compatibility; the kernel never builds the
    (points x monomials) matrix, so there is nothing to batch and results do not depend on it.
    ``devices`` (not in the reference): a list of GPU indices, or ``"all"``, shards the points
    over those GPUs from this one process (``multi.eval_multi_gpu``); default: the current GPU.
    """
    verify_input = verify_input or DEBUG
    exponents = _as_host_int(exponents)
    generating_points = _as_host_float(generating_points)
    _, coefficients, _, _, _, xx = rectify_eval_input(xx, coefficients, exponents, verify_input)
    if devices is not None:
        from .multi import eval_multi_gpu
        return convert_eval_output(eval_multi_gpu(xx, coefficients, exponents, generating_points,
                                                  None if devices == "all" else devices))
    plan = get_plan(exponents, generating_points)
    return convert_eval_output(plan.eval(xx, coefficients))


def deriv_newt_eval(x, coefficients, exponents, generating_points, derivative_order_along):
    """Values of a partial derivative of polynomial(s) in Newton form
    (``polynomials/utils.py:8-112``).  ``x`` (P, m); ``coefficients`` (N,) -> result (P,) as in
    the reference, or (N, k) -> (P, k); ``derivative_order_along`` m non-negative ints."""
    exponents = _as_host_int(exponents)
    generating_points = _as_host_float(generating_points)
    m = exponents.shape[1]
    orders = np.asarray(derivative_order_along).reshape(-1)
    if orders.shape[0] != m:
        raise ValueError(f"derivative_order_along needs {m} entries, {orders.shape[0]} given")
    if not np.issubdtype(orders.dtype, np.integer):
        raise TypeError("derivative orders must be integers")
    if (orders < 0).any():
        raise ValueError("derivative orders must be non-negative")
    if x.ndim != 2 or x.shape[1] != m:
        raise ValueError(f"points of shape (P, {m}) expected, {tuple(x.shape)} given")
    if coefficients.shape[0] != exponents.shape[0]:
        raise ValueError(f"the coefficient amount {coefficients.shape[0]} does not match the "
                         f"amount of monomials {exponents.shape[0]}")
    one_d = coefficients.ndim == 1
    c2 = coefficients.reshape(-1, 1) if one_d else coefficients
    out = get_plan(exponents, generating_points).eval_deriv(x, c2, orders)
    return out.reshape(-1) if one_d else out


def eval_newton_monomials(x, exponents, generating_points, verify_input=False,
                          triangular=False):
    """The (points x monomials) matrix of Newton monomial values."""
    exponents = _as_host_int(exponents)
    generating_points = _as_host_float(generating_points)
    _, m = exponents.shape
    _, x = rectify_query_points(x, m)
    if verify_input:
        _require_dtype("x", x, "float64")
        _require_dtype("exponents", exponents, "int64")
    return get_plan(exponents, generating_points).monomials(x, triangular)


def dds(fct_values, tree):
    """Newton coefficients from function values on the unisolvent nodes, always 2-D (N, k).

    ``tree`` is anything exposing ``multi_index.exponents`` and ``grid.generating_points``:
    the reference's ``MultiIndexTree`` or :class:`minterpy_b200.MultiIndexTree`.
    """
    exponents = np.asarray(tree.multi_index.exponents)
    generating_points = np.asarray(tree.grid.generating_points)
    values = fct_values.reshape(-1, 1) if fct_values.ndim == 1 else fct_values
    plan = get_plan(exponents, generating_points)
    return plan.dds(values)
