"""Route an installed minterpy's hot path through the CUDA library.

``install()`` rebinds the names the reference binds at import time
(``polynomials/newton_polynomial.py:10-11``, ``interpolation.py:15``,
``transformations/utils.py:10,18``) and the operator builders of ``LagrangeToNewton`` /
``NewtonToLagrange`` (``transformations/lagrange.py:23``, ``transformations/newton.py:32``).
The callers either side of the path follow from the same rebinding: Lagrange arithmetic
(``polynomials/lagrange_polynomial.py:29-187``) only composes the patched transformation and
evaluation; the regression matrix (``extras/regression/ordinary_regression.py:8``) and
``NewtonPolynomial.diff`` (``polynomials/newton_polynomial.py:12``) bind their kernels by name
and are rebound here; ``MultiIndexSet.is_complete`` (``core/multi_index.py:20,128``) gets the
linear-time completeness check.  ``uninstall()`` restores the originals.
"""
import importlib

from . import api
from .polynomials import DDSOperator, NodeEvalOperator, is_lexicographically_complete

_saved = []


def _set(obj, name, value):
    _saved.append((obj, name, obj.__dict__.get(name, getattr(obj, name))))
    setattr(obj, name, value)


def install():
    if _saved:
        return
    mods = {n: importlib.import_module(n) for n in (
        "minterpy.utils", "minterpy.dds", "minterpy.interpolation",
        "minterpy.polynomials.newton_polynomial", "minterpy.transformations.utils",
        "minterpy.transformations.lagrange", "minterpy.transformations.newton",
        "minterpy.polynomials.utils", "minterpy.core.multi_index", "minterpy.core.utils",
        "minterpy.extras.regression.ordinary_regression")}
    _set(mods["minterpy.utils"], "eval_newton_polynomials", api.eval_newton_polynomials)
    _set(mods["minterpy.utils"], "eval_newton_monomials", api.eval_newton_monomials)
    _set(mods["minterpy.polynomials.newton_polynomial"], "eval_newton_polynomials",
         api.eval_newton_polynomials)
    for name in ("minterpy.dds", "minterpy.interpolation",
                 "minterpy.polynomials.newton_polynomial", "minterpy.transformations.utils"):
        _set(mods[name], "dds", api.dds)
    _set(mods["minterpy.transformations.utils"], "eval_newton_monomials",
         api.eval_newton_monomials)

    reg = mods["minterpy.extras.regression.ordinary_regression"]
    _set(reg, "eval_newton_polynomials", api.eval_newton_polynomials)
    _set(reg, "eval_newton_monomials", api.eval_newton_monomials)
    _set(mods["minterpy.polynomials.utils"], "deriv_newt_eval", api.deriv_newt_eval)
    _set(mods["minterpy.polynomials.newton_polynomial"], "deriv_newt_eval", api.deriv_newt_eval)
    for name in ("minterpy.core.utils", "minterpy.core.multi_index"):
        _set(mods[name], "is_lexicographically_complete", is_lexicographically_complete)

    # The operators must BE reference operators (tests/test_transformation.py:81 of the reference
    # checks isinstance(operator, OperatorABC)): derive installed variants from both the CUDA
    # operator and OperatorABC; `op @ op` returns the reference's own MatrixOperator.
    operator_abc = importlib.import_module("minterpy.core.ABC").OperatorABC
    ref_matrix_operator = importlib.import_module("minterpy.schemes.matrix_operator").MatrixOperator

    def _installed(base):
        def __init__(self, transformation):
            operator_abc.__init__(self, transformation, None)
            base.__init__(self, transformation.grid, transformation)

        def _make_matrix_operator(self, matrix):
            return ref_matrix_operator(self.transformation, matrix)

        return type(base.__name__, (base, operator_abc),
                    {"__init__": __init__, "_make_matrix_operator": _make_matrix_operator,
                     "__doc__": base.__doc__})

    installed_dds, installed_node_eval = _installed(DDSOperator), _installed(NodeEvalOperator)

    def l2n_operator(transformation):
        grid = transformation.grid
        if grid.multi_index.is_complete and not transformation.origin_poly.indices_are_separate:
            return installed_dds(transformation)
        return _orig_l2n(transformation)

    def n2l_operator(transformation):
        grid = transformation.grid
        if not transformation.origin_poly.indices_are_separate:
            return installed_node_eval(transformation)
        return _orig_n2l(transformation)

    l2n_cls = mods["minterpy.transformations.lagrange"].LagrangeToNewton
    n2l_cls = mods["minterpy.transformations.newton"].NewtonToLagrange
    _orig_l2n = l2n_cls.__dict__["_get_transformation_operator"]
    _orig_n2l = n2l_cls.__dict__["_get_transformation_operator"]
    _set(l2n_cls, "_get_transformation_operator", l2n_operator)
    _set(n2l_cls, "_get_transformation_operator", n2l_operator)


def uninstall():
    while _saved:
        obj, name, value = _saved.pop()
        setattr(obj, name, value)
