"""Light host-side stand-ins for the reference's data model, for use where minterpy itself
is not installed.  Index sets and generating points either come from the reference as arrays
or from ``generators.py`` (``MultiIndexSet.from_degree``, ``Grid(mi)``: bit-identical to the
reference's); the hot path runs in the CUDA library:

* ``NewtonPolynomial.__call__``   ~ ``core/ABC/multivariate_polynomial_abstract.py:94-119``
  -> ``polynomials/newton_polynomial.py:30-54``
* ``LagrangeToNewton`` / ``NewtonToLagrange`` ~ ``transformations/lagrange.py:18-23``,
  ``transformations/newton.py:27-32``, ``core/ABC/transformation_abstract.py:84-180``
* ``Interpolator`` / ``interpolate`` ~ ``interpolation.py:27-181``
"""
import ctypes

import numpy as np

from . import api, generators
from . import _lib
from ._lib import check, lib
from .plan import _is_torch, get_plan


def is_lexicographically_complete(exponents):
    """``core/utils.py:234-262`` for a sorted exponent matrix, through ``mpx_index_is_complete``
    (host code of the library; one merge per dimension instead of the O(N^2 m) search)."""
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    if e.ndim != 2:
        raise ValueError("exponents must be a 2-D array (N x m)")
    flag = ctypes.c_int(0)
    check(lib().mpx_index_is_complete(e.ctypes.data, e.shape[0], e.shape[1], ctypes.byref(flag)))
    return bool(flag.value)


class MultiIndexSet:
    """Holds a lexicographically sorted exponent matrix (N x m, int64)
    (``core/multi_index.py``; only what the hot path reads)."""

    def __init__(self, exponents, lp_degree=None):
        exponents = np.asarray(exponents)
        if exponents.ndim != 2:
            raise ValueError("exponents must be a 2-D array (N x m)")
        self.exponents = np.ascontiguousarray(exponents, dtype=np.int64)
        self.lp_degree = generators.verify_lp_degree(lp_degree)
        self._poly_degree = None
        self._is_complete = None

    @classmethod
    def from_degree(cls, spatial_dimension, poly_degree, lp_degree=None):
        """``core/multi_index.py:70-77``."""
        if type(poly_degree) is not int:
            raise TypeError("only integer polynomial degrees are supported.")
        if type(spatial_dimension) is not int:
            raise TypeError("spatial dimension must be given as integer.")
        lp = generators.verify_lp_degree(lp_degree)
        return cls(generators.get_exponent_matrix(spatial_dimension, poly_degree, lp), lp_degree=lp)

    def __len__(self):
        return self.exponents.shape[0]

    @property
    def spatial_dimension(self):
        return self.exponents.shape[1]

    @property
    def poly_degree(self):
        if self._poly_degree is None:
            self._poly_degree = generators.poly_degree_of(self.exponents, self.lp_degree)
        return self._poly_degree

    @property
    def is_complete(self):
        if self._is_complete is None:
            self._is_complete = is_lexicographically_complete(self.exponents)
        return self._is_complete


class Grid:
    """Index set + generating points; unisolvent nodes as in ``core/grid.py:19-24``.  Without
    generating points the Leja-ordered Chebyshev points of the set's polynomial degree are
    used, as in ``core/grid.py:66-95``."""

    def __init__(self, multi_index, generating_points=None, generating_values=None):
        if not isinstance(multi_index, MultiIndexSet):
            raise TypeError(f"the indices must be given as {MultiIndexSet} class instance")
        self.multi_index = multi_index
        if generating_points is None:
            if generating_values is None:
                generating_values = generators.leja_ordered_chebyshev(multi_index.poly_degree)
            generating_points = generators.generating_points_from_values(
                multi_index.spatial_dimension, generating_values)
        self.generating_values = generating_values
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
            # no grid: the default one of the index set (Leja-ordered Chebyshev points)
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

    def diff(self, order):
        """``polynomials/newton_polynomial.py:68-86``: the derivative's values on the
        unisolvent nodes (its Lagrange coefficients), then divided differences.  Both steps
        run on the device; torch coefficients stay there."""
        order = np.asarray(order).reshape(-1)
        if order.shape[0] != self.spatial_dimension:
            raise ValueError(f"inconsistent number of elements in 'order': "
                             f"{order.shape[0]} given, {self.spatial_dimension} expected")
        nodes = self.grid.unisolvent_nodes
        if _is_torch(self.coeffs):
            import torch
            nodes = torch.as_tensor(nodes, device=self.coeffs.device)
        lag = api.deriv_newt_eval(nodes, self.coeffs, self.multi_index.exponents,
                                  self.grid.generating_points, order)
        newt = api.dds(lag, self.grid.tree)
        if self.coeffs.ndim == 1:
            newt = newt.reshape(-1)
        return NewtonPolynomial(self.multi_index, newt, grid=self.grid)

    def partial_diff(self, dim, order=1):
        """``polynomials/newton_polynomial.py:55-66``."""
        if not 0 <= dim < self.spatial_dimension:
            raise ValueError(f"dimension {dim} is out of range for a {self.spatial_dimension}-"
                             f"dimensional polynomial")
        orders = np.zeros(self.spatial_dimension, dtype=np.int64)
        orders[dim] = order
        return self.diff(orders)


class LagrangePolynomial(_Polynomial):
    """Coefficients are the values on the unisolvent nodes (no direct evaluation, as in the
    reference: ``polynomials/lagrange_polynomial.py``).

    ``+``, ``-`` and ``*`` follow ``polynomials/lagrange_polynomial.py:29-187``: both operands
    go to the Newton basis (dds kernel), a result index set and its default grid are generated
    (sum/difference: the larger degree and lp-degree; product: the SUM of the degrees and of the
    lp-degrees, as the reference does), both Newton polynomials are evaluated on the result's
    unisolvent nodes (evaluation kernel) and the values are combined there.  With torch
    coefficients nothing but the node coordinates crosses the host-device boundary."""

    def _combine(self, other, what):
        if not isinstance(other, LagrangePolynomial):
            return NotImplemented
        if other.spatial_dimension != self.spatial_dimension:
            raise NotImplementedError("operands of different spatial dimension are not supported")
        newt_a = LagrangeToNewton(self)()
        newt_b = LagrangeToNewton(other)()
        mi_a, mi_b = self.multi_index, other.multi_index
        if what == "mul":
            degree = int(mi_a.poly_degree + mi_b.poly_degree)
            lp = mi_a.lp_degree + mi_b.lp_degree
        else:
            degree = int(max(mi_a.poly_degree, mi_b.poly_degree))
            lp = max(mi_a.lp_degree, mi_b.lp_degree)
        res_mi = MultiIndexSet.from_degree(self.spatial_dimension, degree, lp)
        res_grid = Grid(res_mi)
        nodes = res_grid.unisolvent_nodes
        if _is_torch(self.coeffs) or _is_torch(other.coeffs):
            import torch
            ref = self.coeffs if _is_torch(self.coeffs) else other.coeffs
            nodes = torch.as_tensor(nodes, device=ref.device)
        va, vb = newt_a(nodes), newt_b(nodes)
        res = va + vb if what == "add" else va - vb if what == "sub" else va * vb
        return LagrangePolynomial(res_mi, res, grid=res_grid)

    def __add__(self, other):
        return self._combine(other, "add")

    def __sub__(self, other):
        return self._combine(other, "sub")

    def __mul__(self, other):
        return self._combine(other, "mul")


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


def _is_operator(x):
    """True for anything that represents a transformation matrix: the operators below, or an
    ``OperatorABC`` of an installed minterpy (``core/ABC/operator_abstract.py:25``)."""
    return hasattr(x, "array_repr_full") and not hasattr(x, "shape")


class _Operator:
    """What the reference's ``OperatorABC`` gives its operators
    (``core/ABC/operator_abstract.py:25-100``): the cached full matrix, the sub-matrix of the
    active monomials, and ``op @ op`` -> the operator of the matrix product
    (``schemes/matrix_operator.py:13-22``).  ``patch.install()`` derives the installed variants
    from this class AND ``OperatorABC``."""

    transformation = None
    _array_repr_full = None

    def _get_array_repr(self):
        raise NotImplementedError

    def _make_matrix_operator(self, matrix):
        return MatrixOperator(matrix)

    @property
    def array_repr_full(self):
        if self._array_repr_full is None:
            self._array_repr_full = self._get_array_repr()
        return self._array_repr_full

    @property
    def array_repr_sparse(self):
        """``operator_abstract.py:82-100``: the columns of the active monomials when the
        polynomial keeps its indices apart from the grid's, else the full matrix."""
        full = self.array_repr_full
        poly = getattr(self.transformation, "origin_poly", None)
        if poly is not None and getattr(poly, "indices_are_separate", False):
            full = full[:, poly.active_monomials]
        return full

    def _chain(self, other):
        return self._make_matrix_operator(self.array_repr_full @ other.array_repr_full)


class DDSOperator(_Operator):
    """Lagrange -> Newton as an operator: ``op @ coeffs`` runs the dds kernel."""

    def __init__(self, grid, transformation=None):
        self.grid = grid
        self.transformation = transformation

    def __matmul__(self, other):
        if _is_operator(other):
            return self._chain(other)
        other_nd = other.ndim
        out = api.dds(other, self.grid.tree)
        return out.reshape(-1) if other_nd == 1 else out

    def _get_array_repr(self):
        """The N x N matrix (dds of the identity, like ``build_l2n_matrix_dds``)."""
        plan = get_plan(np.asarray(self.grid.multi_index.exponents),
                        np.asarray(self.grid.generating_points))
        return plan.dds(np.eye(len(self.grid.multi_index)), many_zeros=True)

    def full_on_device(self, device=None):
        """The same matrix as a CUDA tensor: identity built and transformed on the device, in
        place, with the kernel instantiation for arrays full of zeros (``MPX_DDS_MANY_ZEROS``)."""
        import torch
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        plan = get_plan(np.asarray(self.grid.multi_index.exponents),
                        np.asarray(self.grid.generating_points), device=dev.index)
        eye = torch.eye(len(self.grid.multi_index), dtype=torch.float64, device=dev)
        return plan.dds(eye, many_zeros=True, inplace=True)


class NodeEvalOperator(_Operator):
    """Newton -> Lagrange as an operator: evaluation at the unisolvent nodes."""

    def __init__(self, grid, transformation=None):
        self.grid = grid
        self.transformation = transformation

    def __matmul__(self, other):
        if _is_operator(other):
            return self._chain(other)
        nodes = self.grid.unisolvent_nodes
        if _is_torch(other):
            import torch
            nodes = torch.as_tensor(nodes, device=other.device)
        e, gp = self.grid.multi_index.exponents, self.grid.generating_points
        c2 = other.reshape(-1, 1) if other.ndim == 1 else other
        out = get_plan(e, gp).eval(nodes, c2)
        return out.reshape(-1) if other.ndim == 1 else out

    def _get_array_repr(self):
        """The N x N matrix of monomial values at the nodes (``transformations/utils.py:35-55``)."""
        return api.eval_newton_monomials(self.grid.unisolvent_nodes,
                                         self.grid.multi_index.exponents,
                                         self.grid.generating_points, triangular=True)

    def full_on_device(self, device=None):
        """The same matrix as a CUDA tensor."""
        import torch
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        return api.eval_newton_monomials(torch.as_tensor(self.grid.unisolvent_nodes, device=dev),
                                         self.grid.multi_index.exponents,
                                         self.grid.generating_points, triangular=True)


class MatrixOperator(_Operator):
    """A dense operator (``schemes/matrix_operator.py:10-26``): ``op @ coeffs`` is one pass of
    ``mpx_apply_dense`` over the matrix, which stays on the device; ``op @ op`` gives the
    operator of the matrix product.  ``lower_triangular=True`` (every Lagrange<->Newton matrix of
    a sorted index set, ``transformations/utils.py:35-55``) lets the kernel skip the zero half."""

    def __init__(self, matrix, lower_triangular=False, transformation=None):
        self.lower_triangular = bool(lower_triangular)
        if self.lower_triangular:
            # the kernel skips what lies right of the diagonal: make that part exactly zero, so
            # that the matrix a caller reads back is the operator that is applied
            matrix = matrix.tril() if _is_torch(matrix) else np.tril(matrix)
        self._host = None if _is_torch(matrix) else np.ascontiguousarray(matrix, dtype=np.float64)
        self._dev = matrix.contiguous() if _is_torch(matrix) else None
        self.transformation = transformation

    def _on_device(self):
        if self._dev is None:
            import torch
            self._dev = torch.as_tensor(self._host, device="cuda")
        return self._dev

    def full_on_device(self, device=None):
        """The matrix as a CUDA tensor (on `device` if given)."""
        T = self._on_device()
        return T if device is None else T.to(device)

    def _get_array_repr(self):
        if self._host is None:
            self._host = self._dev.cpu().numpy()
        return self._host

    def __matmul__(self, other):
        if _is_operator(other):
            return self._chain(other)
        import torch
        T = self._on_device()
        as_numpy = not _is_torch(other)
        c = torch.as_tensor(np.ascontiguousarray(other, dtype=np.float64) if as_numpy else other,
                            dtype=torch.float64, device=T.device)
        if c.shape[0] != T.shape[1]:
            raise ValueError(f"operator of shape {tuple(T.shape)} cannot be applied to "
                             f"{c.shape[0]} coefficients")
        c2 = (c.reshape(-1, 1) if c.ndim == 1 else c).contiguous()
        out = torch.empty((T.shape[0], c2.shape[1]), dtype=torch.float64, device=T.device)
        stream = torch.cuda.current_stream(T.device).cuda_stream
        flags = _lib.MPX_DENSE_LOWER if self.lower_triangular else 0
        check(lib().mpx_apply_dense_ex(T.device.index, T.data_ptr(), T.shape[0], T.shape[1],
                                       T.stride(0), c2.data_ptr(), c2.shape[1], c2.stride(0),
                                       out.data_ptr(), out.stride(0), flags, stream))
        if c.ndim == 1:
            out = out.reshape(-1)
        return out.cpu().numpy() if as_numpy else out


class LagrangeToNewton(_Transformation):
    origin_type = LagrangePolynomial
    target_type = NewtonPolynomial

    def _get_transformation_operator(self):
        """Divided differences where the index set is complete; otherwise, like the reference
        (``transformations/utils.py:155-179``), the dense inverse of the Newton->Lagrange matrix
        (the monomials at the unisolvent nodes), built and inverted on the device."""
        if self.grid.multi_index.is_complete:
            return DDSOperator(self.grid, self)
        import torch
        n2l = NodeEvalOperator(self.grid).full_on_device()
        # the Newton->Lagrange matrix is lower triangular (triangular=True above), hence its
        # inverse: solve against the identity as a triangular system
        eye = torch.eye(n2l.shape[0], dtype=torch.float64, device=n2l.device)
        l2n = torch.linalg.solve_triangular(n2l, eye, upper=False)
        return MatrixOperator(l2n, lower_triangular=True, transformation=self)


class NewtonToLagrange(_Transformation):
    origin_type = NewtonPolynomial
    target_type = LagrangePolynomial

    def _get_transformation_operator(self):
        return NodeEvalOperator(self.grid, self)


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


def interpolate(fct, spatial_dimension, poly_degree=None, lp_degree=None):
    """``interpolation.py:161-181``: ``interpolate(fct, spatial_dimension, poly_degree,
    lp_degree)`` builds the index set and its default grid; ``interpolate(fct, grid)`` takes a
    ready grid (arrays that came from the reference).  Returns the Newton interpolant."""
    if isinstance(spatial_dimension, Grid):
        return Interpolator(spatial_dimension)(fct)
    if poly_degree is None:
        raise TypeError("interpolate(fct, spatial_dimension, poly_degree[, lp_degree])")
    mi = MultiIndexSet.from_degree(spatial_dimension, poly_degree, lp_degree)
    return Interpolator(Grid(mi))(fct)
