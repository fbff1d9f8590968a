import numpy as np

# fixed-step parity bar of BASELINE.json: <= 1e-12 relative in fp64 (stricter
# than the reference's own GPU-vs-CPU bar of rel 1e-11 / abs 1e-13,
# cudapixelsim_t.cpp:46-47)
REL_TOL = 1e-12
ABS_TOL = 1e-14


def max_rel_diff(a, b, abs_floor=ABS_TOL):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    d = np.abs(a - b)
    d = np.where(d <= abs_floor, 0.0, d)
    return float(np.max(d / np.maximum(np.abs(b), 1e-300)))


def assert_parity(a, b, rel=REL_TOL, what=""):
    r = max_rel_diff(a, b)
    assert r <= rel, f"{what}: max relative difference {r:.3e} > {rel:.1e}"
    return r
