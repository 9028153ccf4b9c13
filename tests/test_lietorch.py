"""lietorch_backends (SURVEY 8b third module, 8f N1): this repository's sm_100a / host build against
  * tests/golden/lie_ref.npz -- outputs of the reference's own lietorch_cpu.cpp for every m.def, all four groups (oracle/make_golden_lie.py)
  * the reference's OWN test suite dpvo/lietorch/run_tests.py:16-226 (its only golden tests: fp64 identities and gradient checks at
    atol 1e-8; 1e-3 for Sim3, 1e-6 for matrix / vec), imported unmodified from baseline/_ref and run over this module.
CPU-tensor legs run here; the CUDA legs are the `gpu` tests."""
import ctypes
import importlib
import importlib.util
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "wild-video-3d-reconstruction_b200"
REF_INSTALL = os.path.join(ROOT, "baseline", "_ref")
GROUPS = {1: "SO3", 2: "RxSO3", 3: "SE3", 4: "Sim3"}


@pytest.fixture(scope="module")
def lb():
    pkg = importlib.import_module(PKG)
    pkg.install_as_dpvo_backends()
    mod = importlib.import_module("lietorch_backends")
    assert os.path.dirname(os.path.abspath(mod.__file__)) == os.path.dirname(os.path.abspath(pkg.__file__)), mod.__file__
    return mod


@pytest.fixture(scope="module")
def lie_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "lie_ref.npz"))


def _golden_calls(G, g):
    pat = re.compile(rf"g{g}_call(\d+)_(\w+?)__(\w+?)__out(\d+)$")
    calls = {}
    for k in G.files:
        m = pat.match(k)
        if m:
            calls.setdefault((int(m.group(1)), m.group(2), tuple(m.group(3).split("_"))), {})[int(m.group(4))] = G[k]
    return sorted(calls.items())


def _check_against_golden(lb, G, device, dtype, tol):
    n_checked = 0
    for g in GROUPS:
        for (_, name, args), outs in _golden_calls(G, g):
            ins = [torch.from_numpy(G[f"g{g}_in_{a}"]).to(dtype).to(device) for a in args]
            r = getattr(lb, name)(g, *ins)
            r = [r] if torch.is_tensor(r) else list(r)
            assert len(r) == len(outs), (name, len(r), len(outs))
            for j, t in enumerate(r):
                want = outs[j]
                assert tuple(t.shape) == want.shape and t.dtype == dtype and t.device.type == device, (name, t.shape, want.shape)
                err = np.abs(t.double().cpu().numpy() - want).max() / max(1.0, np.abs(want).max())
                assert err <= tol, f"{GROUPS[g]}.{name}{args} out{j}: {err:.3e} > {tol}"
                n_checked += 1
    assert n_checked >= 4 * 24


def test_matches_reference_cpu_build_fp64_on_cpu_tensors(lb, lie_golden):
    _check_against_golden(lb, lie_golden, "cpu", torch.float64, 1e-12)


def test_matches_reference_cpu_build_fp32_on_cpu_tensors(lb, lie_golden):
    # the reference's float instantiation loses ~1e-5 in the (1 - cos) / theta^2 coefficients at small angles; ours evaluates them in double
    _check_against_golden(lb, lie_golden, "cpu", torch.float32, 2e-5)


def test_argument_checks(lb):
    X = torch.zeros(4, 7, dtype=torch.float64); X[:, 6] = 1
    with pytest.raises(RuntimeError):
        lb.inv(3, X[:, :6].contiguous())                     # wrong width
    with pytest.raises(RuntimeError):
        lb.inv(3, torch.zeros(7, 4, dtype=torch.float64).t())   # not contiguous (lietorch.cpp:7 CHECK_CONTIGUOUS)
    with pytest.raises(RuntimeError):
        lb.mul(3, X, X[:2])                                  # batch mismatch
    with pytest.raises(RuntimeError):
        lb.inv(9, X)                                         # unknown group
    with pytest.raises(RuntimeError):
        lb.inv(3, X.half())
    assert lb.inv(3, X[:0]).shape == (0, 7)                  # empty batch


def _reference_suite():
    """the reference's run_tests.py, unmodified, bound to its own Python layer (dpvo/lietorch/*.py) over OUR lietorch_backends"""
    if not os.path.exists(os.path.join(REF_INSTALL, "dpvo", "lietorch", "run_tests.py")):
        pytest.skip("baseline/_ref not installed (python baseline/install_ref.py in the authoring container)")
    pkg = importlib.import_module(PKG)
    pkg.install_as_dpvo_backends()
    if REF_INSTALL not in sys.path:
        sys.path.insert(0, REF_INSTALL)
    lt = importlib.import_module("dpvo.lietorch")
    import lietorch_backends
    assert os.path.dirname(os.path.abspath(lietorch_backends.__file__)) == os.path.dirname(os.path.abspath(pkg.__file__))
    sys.modules.setdefault("lietorch", lt)
    sys.modules.setdefault("gradcheck", importlib.import_module("dpvo.lietorch.gradcheck"))
    spec = importlib.util.spec_from_file_location("ref_lie_run_tests", os.path.join(REF_INSTALL, "dpvo", "lietorch", "run_tests.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _run_reference_suite(device):
    m = _reference_suite()
    torch.manual_seed(0)
    for G in (m.SO3, m.RxSO3, m.SE3, m.Sim3):           # run_tests.py:229-235
        m.test_exp_log(G, device=device); m.test_inv(G, device=device); m.test_adj(G, device=device); m.test_act(G, device=device)
    for G in (m.SO3, m.RxSO3, m.SE3, m.Sim3):           # run_tests.py:237-252
        tol = 1e-3 if G is m.Sim3 else 1e-8
        m.test_exp_log_grad(G, device=device, tol=tol); m.test_inv_log_grad(G, device=device, tol=tol)
        for fn in (m.test_adj_grad, m.test_adjT_grad, m.test_act_grad, m.test_matrix_grad, m.extract_translation_grad, m.test_vec_grad,
                   m.test_fromvec_grad):
            fn(G, device=device)


def test_reference_own_suite_passes_on_cpu():
    _run_reference_suite("cpu")


@pytest.mark.gpu
def test_matches_reference_cpu_build_on_cuda(lb, lie_golden):
    _check_against_golden(lb, lie_golden, "cuda", torch.float64, 1e-12)
    _check_against_golden(lb, lie_golden, "cuda", torch.float32, 2e-5)


@pytest.mark.gpu
def test_reference_own_suite_passes_on_cuda():
    _run_reference_suite("cuda")


@pytest.mark.gpu
def test_broadcast_strides_equal_the_repeated_form(lb, lie_golden):
    """stride 0 on either operand (one element for the whole batch) == the repeat()-materialised call of broadcasting.py:26-29"""
    lib = importlib.import_module(PKG).capi()
    G = lie_golden
    X = torch.from_numpy(G["g3_in_X"]).float().cuda()              # [n, 7]
    n = 4096
    p = torch.randn(n, 4, device="cuda")
    out = torch.empty(n, 4, device="cuda")
    V = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    x1 = X[5:6].contiguous()
    rc = lib.wv_lie_forward(7, 3, 1, V(x1), ctypes.c_int64(0), V(p), ctypes.c_int64(4), V(out), ctypes.c_int64(n), 0, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(out, lb.act4(3, x1.repeat(n, 1).contiguous(), p))
    Y = torch.from_numpy(G["g3_in_Y"]).float().cuda()
    y1 = Y[2:3].contiguous()
    outm = torch.empty(Y.shape[0], 7, device="cuda")
    assert lib.wv_lie_forward(3, 3, 1, V(X), ctypes.c_int64(7), V(y1), ctypes.c_int64(0), V(outm), ctypes.c_int64(Y.shape[0]), 0, None) == 0
    torch.cuda.synchronize()
    assert torch.equal(outm, lb.mul(3, X, y1.repeat(Y.shape[0], 1).contiguous()))
