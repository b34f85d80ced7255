"""GPU: mkfhe_eval_circuit_dev_static -- the circuit entry point that allocates nothing (scratch in the library's file-scope
`static __device__` array; include/mkfhe_b200.h) -- against the existing mkfhe_eval_circuit_dev, through the C ABI, bit for bit
(integer work: the tolerance of every parity test here is ==), on two calls in a row that share the array.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

from util import golden_names, load_golden

pytestmark = pytest.mark.gpu

# XOR of wires 0 / 1 (levels 0-2), gates mixing the levels, a second netlist for the second call with another shape
NET_A = np.array([(0, 1, 2), (0, 2, 3), (1, 2, 4), (3, 4, 5), (5, 2, 6), (0, 6, 7), (4, 3, 8)], dtype=np.uint32)
NET_B = np.array([(1, 0, 2), (2, 0, 3), (2, 1, 4), (3, 4, 5), (0, 5, 6), (1, 5, 7), (6, 7, 8), (8, 2, 9)], dtype=np.uint32)


def _u32p(a):
    return a.ctypes.data_as(C.c_void_p)


def _run(ctx, entry, wires_t, b_t, n_wires, net):
    """call one of the two circuit entry points of the C ABI directly on device pointers"""
    i1, i2, o = (np.ascontiguousarray(net[:, j]) for j in range(3))
    lv = C.c_uint32(0)
    fn = getattr(ctx._L, entry)
    rc = fn(ctx._h, C.c_void_p(wires_t.data_ptr()), None if b_t is None else C.c_void_p(b_t.data_ptr()), n_wires, _u32p(i1), _u32p(i2),
            _u32p(o), len(net), C.byref(lv))
    return rc, int(lv.value)


def _both_entries_agree(ctx, P, wires, wb, nets):
    import torch

    lwe = P.scheme == 1
    dev = torch.device("cuda", torch.cuda.current_device())
    for net in nets:  # two calls in a row on the static array, each against the pooled entry point on the same inputs
        W = wires.shape[0]
        res = {}
        for entry in ("mkfhe_eval_circuit_dev", "mkfhe_eval_circuit_dev_static"):
            w_t = torch.from_numpy(wires.view(np.int32)).to(dev)
            b_t = torch.from_numpy(wb.view(np.int32)).to(dev) if lwe else None
            torch.cuda.synchronize()
            rc, levels = _run(ctx, entry, w_t, b_t, W, net)
            assert rc == 0, ctx._L.mkfhe_last_error(ctx._h)
            ctx.synchronize()
            res[entry] = (w_t.cpu().numpy().view(np.uint32), b_t.cpu().numpy().view(np.uint32) if lwe else None, levels)
        (w0, b0, l0), (w1, b1, l1) = res["mkfhe_eval_circuit_dev"], res["mkfhe_eval_circuit_dev_static"]
        assert l0 == l1 and l0 > 0
        assert (w0 == w1).all()
        if lwe:
            assert (b0 == b1).all()
        written = np.unique(net[:, 2])
        assert (w1[written] != wires[written]).any()  # the call did write its outputs
        untouched = np.setdiff1d(np.arange(W), written)
        assert (w1[untouched] == wires[untouched]).all()


@pytest.mark.parametrize("name", golden_names())
def test_static_scratch_circuit_equals_pooled_circuit_toy(name):
    from gpu_util import ctx_from_params, load_fixture_keys

    g = load_golden(name)
    p = g["p"]
    ctx = ctx_from_params(p)
    load_fixture_keys(ctx, g)
    c1 = np.asarray(g["g_ct1"], dtype=np.uint32).reshape(-1, p.k, p.n)
    c2 = np.asarray(g["g_ct2"], dtype=np.uint32).reshape(-1, p.k, p.n)
    lwe = p.scheme == 1
    W = 10
    wires = np.zeros((W, p.k, p.n), dtype=np.uint32)
    wb = np.zeros(W, dtype=np.uint32)
    wires[0], wires[1] = c1[0], c2[0]
    if lwe:
        wb[0], wb[1] = int(g["g_b1"][0]), int(g["g_b2"][0])
    _both_entries_agree(ctx, ctx.params, wires, wb, [NET_A, NET_B])
    ctx.close()


@pytest.mark.parametrize("paramset", ["STD100_MKNTRU_LWE", "STD100_MKNTRU"])
def test_static_scratch_circuit_equals_pooled_circuit_full_size(paramset):
    """full-size parameter sets (the specialised gate kernels), uniform random keys and dense random input wires; a wide first
    level (150 gates: the one-group kernel) followed by narrow ones (the party-parallel kernel)"""
    import mkfhe_b200 as mk

    method = mk.MKNTRU_LWE if "LWE" in paramset else mk.MKNTRU
    P = mk.derive_params(mk.PARAMSETS[paramset], method)
    rng = np.random.default_rng(77)
    ell, S, lwe = P.digitsG - 1, (2 if P.variant == 0 else 1), P.scheme == 1
    bk = rng.integers(0, P.Q, size=(P.k * S * P.n + 1) * ell * 2 * P.N, dtype=np.uint32)
    pkey = rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32)
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    if lwe:
        rows = P.k * P.N * P.baseKS * P.dKS
        ctx.LoadKeys(bk, pkey, kskA=rng.integers(0, P.qKS, size=rows * P.n, dtype=np.uint32),
                     kskB=rng.integers(0, P.qKS, size=rows, dtype=np.uint32))
    else:
        ctx.LoadKeys(bk, pkey, ksk=rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32),
                     ctnand=rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32))
    n_in, wide = 151, 150
    W = n_in + wide + 3
    wires = np.zeros((W, P.k, P.n), dtype=np.uint32)
    wires[:n_in] = rng.integers(0, P.q, size=(n_in, P.k, P.n), dtype=np.uint32)
    wb = np.zeros(W, dtype=np.uint32)
    wb[:n_in] = rng.integers(0, P.q, size=n_in, dtype=np.uint32)
    first = [(i, i + 1, n_in + i) for i in range(wide)]
    net1 = np.array(first + [(n_in, n_in + 1, W - 3), (W - 3, n_in + 2, W - 2), (W - 2, 0, W - 1)], dtype=np.uint32)
    net2 = np.array(first[:40] + [(n_in + 3, n_in + 4, W - 1)], dtype=np.uint32)
    _both_entries_agree(ctx, P, wires, wb, [net1, net2])
    ctx.close()


def test_static_scratch_rejects_circuits_beyond_the_array():
    """sizes beyond MKFHE_STATIC_SCRATCH_WORDS are refused with a clear error before anything is launched"""
    import torch

    import mkfhe_b200 as mk
    from mkfhe_b200.lib import load

    L = load()
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_LWE"], mk.MKNTRU_LWE)
    cap = L.mkfhe_static_scratch_words()
    assert cap == 2 << 20
    widest = 1
    while L.mkfhe_circuit_static_words(C.byref(P), widest, widest) <= cap:
        widest += 1
    assert widest > 256  # the stated limit is a useful one at k = 2
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    rng = np.random.default_rng(5)
    ell = P.digitsG - 1
    rows = P.k * P.N * P.baseKS * P.dKS
    ctx.LoadKeys(rng.integers(0, P.Q, size=(P.k * P.n + 1) * ell * 2 * P.N, dtype=np.uint32),
                 rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32),
                 kskA=np.zeros(rows * P.n, dtype=np.uint32), kskB=np.zeros(rows, dtype=np.uint32))
    W = 2 * widest + 1
    dev = torch.device("cuda", torch.cuda.current_device())
    w_t = torch.zeros((W, P.k, P.n), dtype=torch.int32, device=dev)
    b_t = torch.zeros(W, dtype=torch.int32, device=dev)
    net = np.array([(i, i + 1, widest + 1 + i) for i in range(widest)], dtype=np.uint32)  # one level of `widest` gates
    before = ctx.launch_count()
    rc, _ = _run(ctx, "mkfhe_eval_circuit_dev_static", w_t, b_t, W, net)
    assert rc == -1  # MKFHE_EINVAL
    msg = ctx._L.mkfhe_last_error(ctx._h).decode()
    assert "static array" in msg and str(cap) in msg
    assert ctx.launch_count() == before
    torch.cuda.synchronize()
    assert int(w_t.abs().sum().item()) == 0  # nothing ran
    ctx.close()
