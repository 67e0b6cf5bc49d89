"""b200_vec_mdot_noalloc (include/petscb200.h): VecMDot with its scratch in a static device array, for callers that may
not allocate.  No counterpart in the reference; the checker here is the existing entry point b200_vec_mdot_local, which
runs the same kernel on the same grid, so the results agree to the last bit (the tolerance the other VecMDot tests use,
1e-13 * sum|x_i y_i|, is asserted first).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(gpu, P, n, nv, seed, offset=0):
    xh = P.random_vector(n + offset, seed)
    ysh = [P.random_vector(n + offset, seed + 1 + k) for k in range(nv)]
    x = gpu.Vec(n + offset, xh)
    ys = [gpu.Vec(n + offset, v) for v in ysh]
    keep = (x, ys)                                           # the views below do not own the memory
    if offset:
        x, ys = x.offset(offset), [y.offset(offset) for y in ys]
        xh, ysh = xh[offset:], [v[offset:] for v in ysh]
    return keep, x, ys, xh, ysh


def _check(gpu, x, ys, xh, ysh, z):
    nv = len(ys)
    z.set(-7.0)                                              # whatever the last call left must be overwritten
    x.mdot_noalloc(ys, z)
    got = z.get_array()[:nv]                                 # same stream: ordered after the kernel, then synchronised
    want = x.mdot(ys, local=True)
    for k in range(nv):
        assert abs(got[k] - want[k]) <= 1e-13 * np.abs(xh * ysh[k]).sum()
    assert np.array_equal(got, want)                         # same kernel, grid and summation order
    assert np.all(z.get_array()[nv:] == -7.0)                # and nothing written beyond nv results


def test_two_calls_in_a_row_match_the_allocating_entry_point(gpu, P):
    z = gpu.Vec(16)
    # large grid first, then a small one (stale partial sums of the first call stay in the scratch), then unaligned
    for n, nv, seed, off in ((1000003, 16, 900, 0), (50021, 7, 930, 0), (4099, 3, 960, 1), (50021, 7, 930, 0), (1, 1, 990, 0)):
        keep, x, ys, xh, ysh = _case(gpu, P, n, nv, seed, off)
        _check(gpu, x, ys, xh, ysh, z)
        _check(gpu, x, ys, xh, ysh, z)                       # the same call again: identical result
        del keep


def test_launch_count_and_no_host_result(gpu, P):
    """One kernel launch per call, and the call itself does not wait for the device."""
    keep, x, ys, xh, ysh = _case(gpu, P, 50021, 5, 1200)
    z = gpu.Vec(16)
    gpu.synchronize()
    before = gpu.launch_count()
    x.mdot_noalloc(ys, z)
    assert gpu.launch_count() - before == 1
    gpu.synchronize()
    assert np.array_equal(z.get_array()[:5], x.mdot(ys, local=True))


def test_sizes_beyond_the_static_scratch_are_refused(gpu, P):
    keep, x, ys, xh, ysh = _case(gpu, P, 1031, 17, 1100)
    z = gpu.Vec(32)
    z.set(3.0)
    with pytest.raises(gpu.PetscB200Error) as e:
        x.mdot_noalloc(ys, z)
    assert e.value.code == 60 and "exceeds the 16 vectors" in str(e.value)
    tab = (gpu.c_dp * 1)(ys[0].p)
    rc = gpu.lib.b200_vec_mdot_noalloc((1 << 30) + 1, x.p, 1, tab, z.p, C.c_void_p(None))
    assert rc == 60 and b"exceeds the maximum" in gpu.lib.b200_last_error()
    assert gpu.lib.b200_vec_mdot_noalloc(-1, x.p, 1, tab, z.p, C.c_void_p(None)) == 60
    assert gpu.lib.b200_vec_mdot_noalloc(10, x.p, 1, tab, None, C.c_void_p(None)) == 85
    assert gpu.lib.b200_vec_mdot_noalloc(10, x.p, 0, tab, None, C.c_void_p(None)) == 0      # nv = 0: nothing to do
    assert np.all(z.get_array() == 3.0)                      # refused calls wrote nothing
    x.mdot_noalloc(ys[:16], z)                               # and the next good call is unaffected
    assert np.array_equal(z.get_array()[:16], x.mdot(ys[:16], local=True))
