"""Oracle-backed stand-in for minterpy_b200.plan.Plan -- CPU tests of the HOST logic only.

The container that has /root/reference has no GPU and the GPU box has no reference, so the
tests that exercise the Python layers above the C ABI (argument rules, composition of the
kernels into transformations, arithmetic, regression, differentiation) swap the device plan
for this class.  The CUDA path itself is tested by the `-m gpu` tests.
"""
import numpy as np

import oracle


class OraclePlan:
    def __init__(self, exponents, generating_points):
        self.e = np.ascontiguousarray(exponents, dtype=np.int64)
        self.gp = np.ascontiguousarray(generating_points, dtype=np.float64)

    is_complete = True

    def eval(self, xx, coeffs):
        return oracle.eval_newton_polynomials(xx, coeffs, self.e, self.gp)

    def eval_deriv(self, xx, coeffs, orders):
        return oracle.deriv_newt_eval(xx, coeffs, self.e, self.gp, orders)

    def monomials(self, xx, triangular=False):
        return oracle.eval_newton_monomials(xx, self.e, self.gp, triangular)

    def dds(self, values, many_zeros=False, inplace=False):
        return oracle.dds(values, self.e, self.gp)


def fake_get_plan(e, gp, device=None):
    return OraclePlan(e, gp)
