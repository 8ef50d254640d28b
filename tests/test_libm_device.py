"""The path's portable libm (include/vspg_libm.h, include/vspg_rsqrtss.h) gives the same bits on the device and on the host:
the kernels, the oracle and -- through LD_PRELOAD=oracle/libportm.so -- the reference renderer all compute with it, which is
what makes whole paths bit-comparable.  Device values come through the C ABI (vspg_debug_libm), host values from the same
header compiled by g++ into oracle/libportm.so."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def portm():
    L = C.CDLL(os.path.join(ROOT, "oracle", "libportm.so"))
    for name in ("logf", "expf", "sinf", "cosf", "tanf", "asinf", "acosf", "atanf", "vspg_portm_embree_rsqrt"):
        getattr(L, name).restype = C.c_float; getattr(L, name).argtypes = [C.c_float]
    for name in ("powf", "atan2f"):
        getattr(L, name).restype = C.c_float; getattr(L, name).argtypes = [C.c_float, C.c_float]
    L.log.restype = C.c_double; L.log.argtypes = [C.c_double]
    return L


def bits(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = a.view(np.uint32).copy()
    b[np.isnan(a)] = 0x7FC00000
    return b


def inputs(rng, lo, hi, n=20000, log_scale=False, special=()):
    x = np.exp(rng.uniform(np.log(lo), np.log(hi), n)) if log_scale else rng.uniform(lo, hi, n)
    return np.concatenate([x.astype(np.float32), np.float32(list(special))])


CASES = [   # fn id, host name, input sampler
    (0, "logf", lambda r: inputs(r, 1e-30, 1e30, log_scale=True, special=[1.0, 0.0, np.inf, 1e-45, 0.99999994, 1.0000001, -1.0])),
    (1, "expf", lambda r: inputs(r, -100, 90, special=[0.0, -0.0, 88.72, -103.97, np.inf, -np.inf])),
    (3, "sinf", lambda r: np.concatenate([inputs(r, -16, 16), inputs(r, -1e6, 1e6, n=4000), np.float32([0.0, -0.0, 6.2831855, 1e-30])])),
    (4, "cosf", lambda r: np.concatenate([inputs(r, -16, 16), inputs(r, -1e6, 1e6, n=4000), np.float32([0.0, 3.1415927, 1.5707964])])),
    (5, "tanf", lambda r: inputs(r, -1.5, 1.5, special=[0.0, 0.7853982])),
    (6, "asinf", lambda r: inputs(r, -1, 1, special=[1.0, -1.0, 0.0, 1.5, 1e-20])),
    (7, "acosf", lambda r: inputs(r, -1, 1, special=[1.0, -1.0, 0.0, 1.5, 0.99999994])),
    (8, "atanf", lambda r: inputs(r, -1e4, 1e4, special=[0.0, np.inf, -np.inf, 1.0])),
    (10, "vspg_portm_embree_rsqrt", lambda r: inputs(r, 1e-30, 1e30, log_scale=True, special=[1.0, 4.0, 0.25])),
]


@pytest.mark.parametrize("fn,name,gen", CASES, ids=[c[1] for c in CASES])
def test_device_libm_equals_host_libm_bit_for_bit(ctx, portm, fn, name, gen):
    x = gen(np.random.default_rng(fn + 1))
    got = ctx.debug_libm(fn, x)
    f = getattr(portm, name)
    want = np.float32([f(float(v)) for v in x])
    assert np.array_equal(bits(got), bits(want)), f"{name}: {(bits(got) != bits(want)).sum()} of {len(x)} differ"


def test_device_powf_atan2f_and_double_log_equal_the_host(ctx, portm):
    rng = np.random.default_rng(9)
    a = np.concatenate([np.exp(rng.uniform(-20, 20, 20000)), [0.0, 1.0, 2.0, 0.5, 10.0]]).astype(np.float32)
    b = np.concatenate([rng.uniform(-8, 8, 20000), [0.0, 3.0, 0.5, -1.0, 2.4]]).astype(np.float32)
    got = ctx.debug_libm(2, a, b)
    want = np.float32([portm.powf(float(u), float(v)) for u, v in zip(a, b)])
    assert np.array_equal(bits(got), bits(want))
    y = np.concatenate([rng.uniform(-5, 5, 20000), [0.0, 0.0, 1.0, -1.0, -0.0]]).astype(np.float32)
    x = np.concatenate([rng.uniform(-5, 5, 20000), [1.0, -1.0, 0.0, 0.0, -1.0]]).astype(np.float32)
    got = ctx.debug_libm(9, y, x)
    want = np.float32([portm.atan2f(float(u), float(v)) for u, v in zip(y, x)])
    assert np.array_equal(bits(got), bits(want))
    z = np.exp(rng.uniform(-30, 30, 20000)).astype(np.float32)
    got = ctx.debug_libm(11, z)
    want = np.float32([np.float32(portm.log(float(v))) for v in z])
    assert np.array_equal(bits(got), bits(want))
