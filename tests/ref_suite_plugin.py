"""pytest plugin for tests/test_reference_suite.py: import the unmodified reference through the
numpy-2 shim, then `minterpy_b200.install()` so that the reference's OWN test files exercise
the CUDA path.  MPX_REF_SUITE_PLAN=oracle swaps the device plan for the oracle-backed stand-in
(CPU run: wiring, shapes, errors); =cuda leaves the CUDA library in place (GPU run)."""
import os


def _setup():
    # runs at plugin import (-p): before pytest loads the reference's conftest.py
    from ref_shim import load_reference
    if load_reference() is None:
        raise RuntimeError("reference not found (baseline/_ref or /root/reference)")
    import minterpy_b200
    from minterpy_b200 import api, polynomials
    mode = os.environ.get("MPX_REF_SUITE_PLAN", "cuda")
    if mode == "unpatched":          # the reference as it is, for comparison
        return
    if mode == "oracle":
        from oracle_plan import fake_get_plan
        api.get_plan = fake_get_plan
        polynomials.get_plan = fake_get_plan
    elif mode != "cuda":
        raise RuntimeError(f"MPX_REF_SUITE_PLAN={mode!r}")
    minterpy_b200.install()


_setup()
