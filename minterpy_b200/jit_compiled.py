"""Alias module: BASELINE.json speaks of ``minterpy.jit_compiled``'s
``eval_newton_polynomials(xx, coeffs, exponents, generating_points)`` and ``dds``; in the
reference snapshot they live in ``minterpy.utils`` / ``minterpy.dds``."""
from .api import dds, eval_newton_monomials, eval_newton_polynomials  # noqa: F401
