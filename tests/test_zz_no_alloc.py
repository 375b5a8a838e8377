"""-m gpu: ink_check_finite_noalloc, the one entry point of libinkb200.so that allocates no device memory.

Who it is for: a host that cannot create an ink_ctx because ink_create cudaMallocs the simulation state — an application
that gave all of HBM to its own arena at start-up (cudaMalloc fails there), or a frame loop in which cudaMalloc / cudaFree
would stall every stream of the process.  Such a host holds the grid and the particles in its own device buffers and lets the
library inspect them in place; the kernel's scratch is a fixed array inside the library image, shared by all calls.

What is checked: through the C ABI, on the same state, its counts equal ink_check_finite's (integers: exact, the bar of every
other integer comparison in this suite), on two calls in a row with different inputs and sizes, so that anything the first
call left in the shared scratch array would show up in the second; fp64 and the N x 3 particle layout; sizes beyond the
stated maximum are refused with a message.  The torch tensors below are the CALLER's buffers (the test plays the host that
owns its memory); the entry point itself is handed plain pointers.

This file sorts last on purpose (pytest -x stops at the first failure: a failure here hides nothing behind it) and passes both
in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


def _noalloc_counts(L, torch, triples, n, stride, dtype, check_norm, out, stream=None):
    """One call through the C ABI; `out` is a 2-element int64 CUDA tensor the entry point overwrites."""
    x, y, z = triples
    rc = L.ink_check_finite_noalloc(C.c_void_p(x), C.c_void_p(y), C.c_void_p(z), n, stride, dtype, check_norm,
                                    C.c_void_p(out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, L.ink_last_error(None)
    torch.cuda.synchronize()
    return [int(v) for v in out.cpu()]


def _state_counts(L, torch, fields, pos, vel, bits, out):
    """checkNanAndInf over a whole state = grid velocity (norm test), particle positions (NaN / Inf only), particle
    velocities (norm test): three calls on the null stream, summed — what ink_check_finite does in one."""
    esz = bits // 8
    tot = np.zeros(2, dtype=np.int64)
    u, v, w = fields
    tot += _noalloc_counts(L, torch, (u.data_ptr(), v.data_ptr(), w.data_ptr()), u.numel(), 1, bits, 1, out)
    for arr, norm in ((pos, 0), (vel, 1)):      # N x 3 arrays: components of one element are `esz` bytes apart, elements 3 apart
        p = arr.data_ptr()
        tot += _noalloc_counts(L, torch, (p, p + esz, p + 2 * esz), arr.shape[0], 3, bits, norm, out)
    return int(tot[0]), int(tot[1])


@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_noalloc_check_finite_equals_ink_check_finite_twice_in_a_row(dtype):
    import torch

    from ink_simulation_b200 import System, _capi

    L = _capi.load()
    bits = 32 if dtype == np.float32 else 64
    out = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    rng = np.random.default_rng(5)
    # two different states, the second larger than the first (more blocks than the first call used) and with other counts
    cases = []
    for dims, npart, seed in ((util.C1, 20000, 1), (util.C2, 70001, 2)):
        u, v, w = util.rand_field(dims, 50 + seed, scale=0.5, dtype=dtype)
        pos = util.box_particles(dims, npart, 60 + seed, dtype=dtype)
        vel = (rng.standard_normal(pos.shape) * 0.1).astype(dtype)
        for k in range(3 + 4 * seed):            # violations of every kind the reference asserts on
            i, j, kk = (int(rng.integers(0, d)) for d in dims)
            (u, v, w)[k % 3][i, j, kk] = (np.nan, np.inf, -np.inf, 3e4)[k % 4]
        for k in range(5 * seed):
            pos[int(rng.integers(0, npart)), k % 3] = (np.nan, np.inf)[k % 2]
            vel[int(rng.integers(0, npart)), k % 3] = (10000.0, -np.inf, np.nan)[k % 3]   # exactly 10000 is not < 10000
        cases.append((dims, (u, v, w), pos, vel))
    results = []
    for dims, fields, pos, vel in cases:
        S = System(*dims, dtype=dtype)
        S.set_velocity(*fields)
        S.set_particles(pos, vel)
        want = S.check_finite()                   # the existing entry point (ink_check_finite), on its own allocated state
        S.close()
        assert want[0] > 0 and want[1] > 0
        dev_fields = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in fields]
        dpos, dvel = torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda()
        got = _state_counts(L, torch, dev_fields, dpos, dvel, bits, out)
        assert got == want, (dims, got, want)
        results.append(got)
    assert results[0] != results[1]
    # and once more back on the first, smaller state: the slots the larger call filled must not leak into it
    dims, fields, pos, vel = cases[0]
    dev_fields = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in fields]
    got = _state_counts(L, torch, dev_fields, torch.from_numpy(pos).cuda(), torch.from_numpy(vel).cuda(), bits, out)
    assert got == results[0]


def test_gpu_noalloc_runs_on_the_callers_stream_and_many_blocks():
    """A range of several hundred blocks on a caller-owned stream, twice back to back without a host sync in between, into two
    different outputs: the second launch must find the ticket cleared by the first."""
    import torch

    from ink_simulation_b200 import _capi

    L = _capi.load()
    n = 3_000_001
    g = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randn(3, n, device="cuda", generator=g)
    b = a.clone()
    bad_a = torch.randint(0, n, (1000,), device="cuda", generator=g).unique()
    big_b = torch.randint(0, n, (777,), device="cuda", generator=g).unique()
    a[1, bad_a] = float("nan")
    b[2, big_b] = 1.0e5
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    for t, o in ((a, 0), (b, 2)):
        rc = L.ink_check_finite_noalloc(C.c_void_p(t[0].data_ptr()), C.c_void_p(t[1].data_ptr()), C.c_void_p(t[2].data_ptr()), n, 1, 32, 1,
                                        C.c_void_p(out.data_ptr() + 8 * o), C.c_void_p(st.cuda_stream))
        assert rc == 0, L.ink_last_error(None)
    st.synchronize()
    # a NaN component also fails `norm() < 10000`, as in the reference
    assert [int(v) for v in out.cpu()] == [bad_a.numel(), bad_a.numel(), 0, big_b.numel()]


def test_gpu_noalloc_refuses_what_it_cannot_cover():
    import torch

    from ink_simulation_b200 import _capi

    L = _capi.load()
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    x = torch.zeros(16, device="cuda")
    p, o = C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr())
    assert L.ink_check_finite_noalloc(p, p, p, _capi.NOALLOC_MAX_N + 1, 1, 32, 1, o, None) == 2
    assert b"INK_NOALLOC_MAX_N" in L.ink_last_error(None)
    assert L.ink_check_finite_noalloc(p, p, p, -1, 1, 32, 1, o, None) == 2
    assert L.ink_check_finite_noalloc(p, p, p, 16, 0, 32, 1, o, None) == 2
    assert L.ink_check_finite_noalloc(p, p, p, 16, 1, 16, 1, o, None) == 2
    assert L.ink_check_finite_noalloc(p, p, p, 16, 1, 32, 1, None, None) == 2
    assert L.ink_check_finite_noalloc(None, p, p, 16, 1, 32, 1, o, None) == 2
    # an empty range is fine and reports zeros
    out.fill_(-1)
    assert L.ink_check_finite_noalloc(None, None, None, 0, 1, 32, 1, o, None) == 0
    torch.cuda.synchronize()
    assert [int(v) for v in out.cpu()] == [0, 0]
