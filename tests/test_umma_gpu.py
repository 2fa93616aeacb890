"""tcgen05 building blocks (csrc/umma.cuh) against NumPy: 16-bit operands (fp16 by default, bf16 builds), fp32 accumulate.
Covers the operand views every tensor-core kernel of the library relies on."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def bf16_round(a):
    """Round to the library's 16-bit operand type."""
    from gnn_uds_b200 import _lib
    if _lib.load().gnnuds_tc_operand_is_half():
        return np.asarray(a, np.float32).astype(np.float16).astype(np.float32)
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


@pytest.fixture(scope="module")
def em():
    e, _ = cases.setup_product("c3_ast_act_5")
    return e


def run(em, mode, act, mat, shift=0, swap=0):
    Ra, Ca = act.shape
    Rm, Km = mat.shape
    nout = Rm if mode == 0 else Ca
    out = np.full((128, nout), np.nan, np.float32)
    act = np.ascontiguousarray(act, np.float32)
    mat = np.ascontiguousarray(mat, np.float32)
    rc = em._lib.gnnuds_debug_umma_selftest(em._h, mode, act.ctypes.data_as(C.c_void_p), Ra, Ca,
                                            mat.ctypes.data_as(C.c_void_p), Rm, Km, shift, swap,
                                            out.ctypes.data_as(C.c_void_p))
    assert rc == 0, em._lib.gnnuds_last_error(em._h)
    return out


@pytest.mark.parametrize("Ca,Nout,shift,Ra", [(64, 64, 0, 128), (192, 128, 0, 136), (128, 64, 5, 144), (384, 64, 8, 200),
                                                (128, 192, 3, 131 + 8)])
def test_kmajor_activations_times_weight_image(em, Ca, Nout, shift, Ra):
    rng = np.random.default_rng(Ca + Nout + shift)
    act = rng.standard_normal((Ra, Ca)).astype(np.float32)
    mat = rng.standard_normal((Nout, Ca)).astype(np.float32)
    got = run(em, 0, act, mat, shift)
    ref = bf16_round(act[shift:shift + 128]).astype(np.float64) @ bf16_round(mat).astype(np.float64).T
    err = np.abs(got - ref).max() / np.abs(ref).max()
    print("mode0 Ca=%d Nout=%d shift=%d err %.3g" % (Ca, Nout, shift, err))
    assert err < 1e-5


@pytest.mark.parametrize("Km,Ca,Ra", [(128, 64, 128), (144, 64, 144), (144, 128, 150), (128, 128, 128)])
def test_weight_image_times_mnmajor_activations(em, Km, Ca, Ra):
    rng = np.random.default_rng(Km + Ca)
    act = rng.standard_normal((Ra, Ca)).astype(np.float32)
    mat = rng.standard_normal((128, Km)).astype(np.float32)
    got = run(em, 1, act, mat)
    ref = bf16_round(mat).astype(np.float64) @ bf16_round(act[:Km]).astype(np.float64)
    err = np.abs(got - ref).max() / np.abs(ref).max()
    print("mode1 Km=%d Ca=%d err %.3g" % (Km, Ca, err))
    assert err < 1e-5
