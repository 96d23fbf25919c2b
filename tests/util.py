"""helpers shared by the parity tests"""
import numpy as np

FP64_RTOL = 1e-12     # north_star: fp64 sums / averages / windowed means, order-of-summation only
FP32_RTOL = 1e-6      # ratios(int32 | float) are float32 cells in the reference (types.h:199-204)


def assert_same(got, want, what="", scale=0.0):
    """bit-exact for integer / index cells; floating point within rtol * max(|got|, |want|, scale).

    `scale` is the magnitude of the INPUT values for results that are differences of rounded
    quantities (windowed means: the reference's recurrence ret[i-1] + (a[i]-a[i-w])/w leaves
    rounding noise of order eps*|a| where the exact answer is 0), so that "order of summation only"
    differences are measured against what was summed, not against a cancelled result."""
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    assert got.dtype == want.dtype, f"{what}: dtype {got.dtype} != {want.dtype}"
    if got.dtype.kind in "iu":
        bad = np.flatnonzero((got != want).reshape(len(got), -1).any(axis=1)) if got.size else []
        assert len(bad) == 0, f"{what}: {len(bad)} cells differ, first at {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
    else:
        rtol = FP64_RTOL if got.dtype == np.float64 else FP32_RTOL
        both_nan = np.isnan(got) & np.isnan(want)
        same_inf = np.isinf(want) & (got == want)
        with np.errstate(invalid="ignore"):
            ok = both_nan | same_inf | (np.abs(got - want) <= rtol * np.maximum(np.maximum(np.abs(want), np.abs(got)), scale))
        bad = np.flatnonzero(~ok)
        assert len(bad) == 0, f"{what}: {len(bad)} cells beyond rtol={rtol}, first at {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
