"""gnnuds_objective_static (include/gnnuds.h): the astlingen objective through the entry point that allocates nothing -
its scratch is a file-scope static __device__ array - against the existing gnnuds_objective, on two calls in a row.
No counterpart in the reference.  This file is collected last on purpose:
a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

import cases
from gnn_uds_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 4 << 20     # GNNUDS_STATIC_SCRATCH_BYTES


class _Dev:
    """Device buffers through the C-ABI (gnnuds_alloc_device / gnnuds_memcpy): the TEST allocates, the entry point does not."""

    def __init__(self, em):
        self.em, self.lib, self.ptrs = em, em._lib, []

    def put(self, arr):
        a = np.ascontiguousarray(arr, np.float32)
        p = C.c_void_p()
        _lib.check(self.em._h, self.lib.gnnuds_alloc_device(self.em._h, C.byref(p), a.nbytes))
        _lib.check(self.em._h, self.lib.gnnuds_memcpy(self.em._h, p, a.ctypes.data_as(C.c_void_p), a.nbytes, 1))
        self.ptrs.append(p)
        return p

    def empty(self, n):
        return self.put(np.full(n, np.nan, np.float32))

    def get(self, p, n):
        out = np.empty(n, np.float32)
        _lib.check(self.em._h, self.lib.gnnuds_memcpy(self.em._h, out.ctypes.data_as(C.c_void_p), p, out.nbytes, 2))
        return out

    def free(self):
        for p in self.ptrs:
            self.lib.gnnuds_free_device(self.em._h, p)
        self.ptrs = []


def _weights(n):
    rng = np.random.default_rng(11)
    fw = np.where(rng.random(n) < 0.4, rng.integers(1, 3, n), 0).astype(np.float32)
    ww = np.zeros(n, np.float32)
    ww[n - 1] = -0.1
    rw = np.where(rng.random(n) < 0.3, 0.01, 0.0).astype(np.float32)
    return fw, ww, rw


def test_objective_static_matches_objective_twice():
    em, inp = cases.setup_product("c3_ast_act_5")
    y1, _ = em.predict(*inp)
    B, T, N = y1.shape[0], y1.shape[1], y1.shape[2]
    q1 = np.ascontiguousarray(np.asarray(inp[0])[:, -1, :, 1], np.float32)
    # second call: another population (more candidates, other values), so that anything left in the shared scratch by the
    # first call would show
    rng = np.random.default_rng(5)
    reps = 7
    y2 = np.concatenate([y1 * rng.uniform(0.5, 1.5, (1, 1, N, 1)).astype(np.float32) for _ in range(reps)])[::-1]
    q2 = np.concatenate([q1] * reps)[::-1] * np.float32(0.75)
    fw, ww, rw = _weights(N)
    gamma = (0.95 ** np.arange(T)).astype(np.float32)
    dev = _Dev(em)
    try:
        dfw, dww, drw, dg = dev.put(fw), dev.put(ww), dev.put(rw), dev.put(gamma)
        for k, (y, q, rough, gm) in enumerate(((y1, q1, drw, dg), (y2, q2, None, None), (y2, q2, drw, dg))):
            nb = y.shape[0]
            dy, dq, d_ref, d_new = dev.put(y), dev.put(q), dev.empty(nb), dev.empty(nb)
            _lib.check(em._h, em._lib.gnnuds_objective(em._h, dy, dq, nb, T, dfw, dww, rough, gm, d_ref, _lib.MEM_DEVICE))
            _lib.check(em._h, em._lib.gnnuds_objective_static(em._h, dy, dq, nb, T, dfw, dww, rough, gm, d_new, 1))
            ref, new = dev.get(d_ref, nb), dev.get(d_new, nb)
            assert np.isfinite(ref).all() and np.isfinite(new).all()
            # the tolerance of test_objective_matches_oracle (tests/test_parity_gpu.py); the two kernels add the same
            # terms in the same order, so the results are in fact bitwise equal
            assert np.allclose(new, ref, rtol=1e-5, atol=1e-5), (k, np.abs(new - ref).max())
            assert np.array_equal(new, ref), (k, np.abs(new - ref).max())
        # enqueue-only form (sync = 0): gnnuds_synchronize waits
        nb = y1.shape[0]
        dy, dq, d_new = dev.put(y1), dev.put(q1), dev.empty(nb)
        _lib.check(em._h, em._lib.gnnuds_objective_static(em._h, dy, dq, nb, T, dfw, dww, drw, dg, d_new, 0))
        _lib.check(em._h, em._lib.gnnuds_synchronize(em._h))
        host = em.objective(y1, inp[0], fw, ww, rw, gamma)
        assert np.array_equal(dev.get(d_new, nb), host)
    finally:
        dev.free()


def test_objective_static_rejects_sizes_beyond_the_static_array():
    em, inp = cases.setup_product("c3_ast_act_5")
    N = em.n_node
    b_max = STATIC_SCRATCH_BYTES // 4 // (128 + N)
    dev = _Dev(em)
    try:
        d = dev.put(np.zeros(64, np.float32))      # never dereferenced: the size check comes first
        rc = em._lib.gnnuds_objective_static(em._h, d, d, b_max + 1, em.seq_out, d, d, None, None, d, 1)
        assert rc == -1
        msg = em._lib.gnnuds_last_error(em._h)
        msg = msg.decode() if isinstance(msg, bytes) else str(msg)
        assert "static" in msg and str(b_max + 1) in msg, msg
        assert em._lib.gnnuds_objective_static(em._h, None, d, 1, em.seq_out, d, d, None, None, d, 1) == -1
    finally:
        dev.free()
