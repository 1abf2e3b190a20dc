"""Light host-side stand-ins for the reference's data model, for use where minterpy itself
is not installed.  They only carry arrays -- index sets and generating points are produced
by the reference (``MultiIndexSet.from_degree`` / ``Grid``) and handed over as arrays -- and
route the hot path through the CUDA library:

* ``NewtonPolynomial.__call__``   ~ ``core/ABC/multivariate_polynomial_abstract.py:94-119``
  -> ``polynomials/newton_polynomial.py:30-54``
* ``LagrangeToNewton`` / ``NewtonToLagrange`` ~ ``transformations/lagrange.py:18-23``,
  ``transformations/newton.py:27-32``, ``core/ABC/transformation_abstract.py:84-180``
* ``Interpolator`` / ``interpolate`` ~ ``interpolation.py:27-181``
"""
import numpy as np

from . import api
from .plan import _is_torch, get_plan


class MultiIndexSet:
    """Holds a lexicographically sorted exponent matrix (N x m, int64)."""

    def __init__(self, exponents, lp_degree=None):
        exponents = np.asarray(exponents)
        if exponents.ndim != 2:
            raise ValueError("exponents must be a 2-D array (N x m)")
        self.exponents = np.ascontiguousarray(exponents, dtype=np.int64)
        self.lp_degree = lp_degree

    def __len__(self):
        return self.exponents.shape[0]

    @property
    def spatial_dimension(self):
        return self.exponents.shape[1]


class Grid:
    """Index set + generating points; unisolvent nodes as in ``core/grid.py:19-24``."""

    def __init__(self, multi_index, generating_points):
        self.multi_index = multi_index
        gp = np.ascontiguousarray(generating_points, dtype=np.float64)
        if gp.ndim != 2 or gp.shape[1] != multi_index.spatial_dimension:
            raise ValueError("generating_points must be (n+1) x m")
        self.generating_points = gp
        self._nodes = None
        self._tree = None

    @property
    def unisolvent_nodes(self):
        if self._nodes is None:
            e = self.multi_index.exponents
            self._nodes = np.take_along_axis(self.generating_points, e, axis=0)
        return self._nodes

    @property
    def tree(self):
        if self._tree is None:
            self._tree = MultiIndexTree(self)
        return self._tree


class MultiIndexTree:
    """What ``dds`` needs from ``core/tree.py``: the grid, checked for completeness."""

    def __init__(self, grid):
        self.grid = grid
        plan = get_plan(grid.multi_index.exponents, grid.generating_points)
        if not plan.is_complete:
            raise ValueError(
                "trying to use the divided difference scheme (multi index tree) with incomplete "
                "multi indices, but DDS only works for complete multi indices (without 'holes').")

    @property
    def multi_index(self):
        return self.grid.multi_index


class _Polynomial:
    def __init__(self, multi_index, coeffs, grid=None, generating_points=None):
        self.multi_index = multi_index
        if grid is None:
            if generating_points is None:
                raise ValueError("either grid or generating_points is required")
            grid = Grid(multi_index, generating_points)
        self.grid = grid
        if coeffs is not None and coeffs.shape[0] != len(multi_index):
            raise ValueError(f"the coefficient amount {coeffs.shape[0]} does not match the "
                             f"amount of monomials {len(multi_index)}")
        self.coeffs = coeffs

    @property
    def unisolvent_nodes(self):
        return self.grid.unisolvent_nodes

    @property
    def spatial_dimension(self):
        return self.multi_index.spatial_dimension


class NewtonPolynomial(_Polynomial):
    def __call__(self, x):
        return api.eval_newton_polynomials(x, self.coeffs, self.multi_index.exponents,
                                           self.grid.generating_points, verify_input=api.DEBUG)


class LagrangePolynomial(_Polynomial):
    """Coefficients are the values on the unisolvent nodes (no direct evaluation, as in the
    reference: ``polynomials/lagrange_polynomial.py``)."""


class _Transformation:
    origin_type = None
    target_type = None

    def __init__(self, origin_poly):
        if type(origin_poly) is not self.origin_type:
            raise TypeError(f"<{type(origin_poly)}> is not <{self.origin_type}>")
        self.origin_poly = origin_poly
        self._op = None

    @property
    def grid(self):
        return self.origin_poly.grid

    @property
    def transformation_operator(self):
        if self._op is None:
            self._op = self._get_transformation_operator()
        return self._op

    def __call__(self, origin_poly=None):
        if origin_poly is None:
            origin_poly = self.origin_poly
        elif type(origin_poly) is not self.origin_type:
            raise TypeError(f"Input polynomial type <{type(origin_poly)}> differs from expected "
                            f"polynomial type <{self.origin_type}>")
        coeffs = self.transformation_operator @ origin_poly.coeffs
        return self.target_type(origin_poly.multi_index, coeffs, grid=origin_poly.grid)


class DDSOperator:
    """Lagrange -> Newton as an operator: ``op @ coeffs`` runs the dds kernel."""

    def __init__(self, grid):
        self.grid = grid
        self._full = None

    def __matmul__(self, other):
        other_nd = other.ndim
        out = api.dds(other, self.grid.tree)
        return out.reshape(-1) if other_nd == 1 else out

    @property
    def array_repr_full(self):
        """The N x N matrix (dds of the identity, like ``build_l2n_matrix_dds``)."""
        if self._full is None:
            self._full = api.dds(np.eye(len(self.grid.multi_index)), self.grid.tree)
        return self._full


class NodeEvalOperator:
    """Newton -> Lagrange as an operator: evaluation at the unisolvent nodes."""

    def __init__(self, grid):
        self.grid = grid
        self._full = None

    def __matmul__(self, other):
        nodes = self.grid.unisolvent_nodes
        if _is_torch(other):
            import torch
            nodes = torch.as_tensor(nodes, device=other.device)
        e, gp = self.grid.multi_index.exponents, self.grid.generating_points
        c2 = other.reshape(-1, 1) if other.ndim == 1 else other
        out = get_plan(e, gp).eval(nodes, c2)
        return out.reshape(-1) if other.ndim == 1 else out

    @property
    def array_repr_full(self):
        """The N x N matrix of monomial values at the nodes (``transformations/utils.py:35-55``)."""
        if self._full is None:
            self._full = api.eval_newton_monomials(self.grid.unisolvent_nodes,
                                                   self.grid.multi_index.exponents,
                                                   self.grid.generating_points, triangular=True)
        return self._full


class LagrangeToNewton(_Transformation):
    origin_type = LagrangePolynomial
    target_type = NewtonPolynomial

    def _get_transformation_operator(self):
        return DDSOperator(self.grid)


class NewtonToLagrange(_Transformation):
    origin_type = NewtonPolynomial
    target_type = LagrangePolynomial

    def _get_transformation_operator(self):
        return NodeEvalOperator(self.grid)


class InterpolationError(Exception):
    pass


class Interpolator:
    """``interpolation.py:27-80`` on a given grid."""

    def __init__(self, grid):
        self.grid = grid
        self.multi_index = grid.multi_index

    def __call__(self, fct):
        try:
            fct_values = fct(self.grid.unisolvent_nodes)
            coeffs = api.dds(fct_values, self.grid.tree).reshape(-1)
        except Exception as e:
            raise InterpolationError(e) from e
        return NewtonPolynomial(self.multi_index, coeffs, grid=self.grid)


def interpolate(fct, grid):
    """``interpolation.py:161-181``; the grid (index set + generating points) is given."""
    return Interpolator(grid)(fct)
