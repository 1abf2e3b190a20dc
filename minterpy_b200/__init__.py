"""minterpy_b200 -- B200-native (sm_100a) Newton evaluation and divided differences for
minterpy's hot path.  See DESIGN.md / INTEGRATION.md."""
from .api import dds, deriv_newt_eval, eval_newton_monomials, eval_newton_polynomials  # noqa: F401
from .multi import eval_multi_gpu  # noqa: F401
from .plan import Plan, clear_plan_cache, get_plan  # noqa: F401
from .polynomials import (Grid, InterpolationError, Interpolator, LagrangePolynomial,  # noqa: F401
                          LagrangeToNewton, MultiIndexSet, MultiIndexTree, NewtonPolynomial,
                          NewtonToLagrange, interpolate)
from .patch import install, uninstall  # noqa: F401
from .regression import OrdinaryRegression, compute_regression_matrix  # noqa: F401

__version__ = "0.1.0"
