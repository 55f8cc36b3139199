"""``cb2_bn_forward_static``: the batch-norm forward for callers that cannot allocate.

Some embedders of the C ABI have no way to hand the kernel a scratch buffer: a plug-in living in a
process whose whole HBM was claimed by the host's own pool, a non-torch host with no caching
allocator, a loop that must never trigger the device-wide synchronisation of ``cudaMalloc``.  For
them the scratch is a fixed 128 KiB array inside the library (limit: C <= 2048).

This test drives the entry point through the C ABI only (ctypes on the shared library, raw
pointers) and checks it against ``cb2_bn_forward`` fed with caller-provided scratch - twice in a
row with different shapes, so the second call meets the first call's leftovers in the shared
array.  The file name sorts last on purpose: a ``pytest -x`` run has executed everything else
before it gets here.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

_LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cloud_b200", "_lib",
                    "libcloud_b200_kernels.so")


def _abi():
    lib = ctypes.CDLL(_LIB)
    vp, i, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    lib.cb2_last_error.restype = ctypes.c_char_p
    lib.cb2_bn_scratch_floats.restype = i
    lib.cb2_bn_scratch_floats.argtypes = [i]
    lib.cb2_bn_static_scratch_max_channels.restype = i
    lib.cb2_bn_forward.restype = i
    lib.cb2_bn_forward.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, f, i, i, vp]
    lib.cb2_bn_forward_static.restype = i
    lib.cb2_bn_forward_static.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, f, f, i, i, vp]
    return lib


def _case(lib, N, C, HW, relu, with_res, seed):
    dev = torch.device("cuda:0")
    torch.manual_seed(seed)
    M = N * HW * HW
    x = (torch.randn(M, C, device=dev) * 1.5 + 0.3).to(torch.bfloat16)
    res = torch.randn(M, C, device=dev).to(torch.bfloat16) if with_res else None
    gamma = torch.rand(C, device=dev) + 0.5
    beta = torch.randn(C, device=dev) * 0.1
    stream = torch.cuda.current_stream().cuda_stream

    def run(static):
        y = torch.empty_like(x)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        save = torch.empty(2 * C, device=dev)
        common = (gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), save.data_ptr(),
                  save[C:].data_ptr(), None, M, C, 1e-5, 0.1, int(relu), 1, stream)
        rptr = None if res is None else res.data_ptr()
        if static:
            rc = lib.cb2_bn_forward_static(x.data_ptr(), rptr, y.data_ptr(), *common)
        else:
            sums = torch.zeros(lib.cb2_bn_scratch_floats(C), device=dev)      # caller-provided scratch
            rc = lib.cb2_bn_forward(x.data_ptr(), rptr, y.data_ptr(), sums.data_ptr(), *common)
        assert rc == 0, lib.cb2_last_error().decode()
        torch.cuda.synchronize()
        return y.float(), rm.clone(), rv.clone(), save.clone()

    want = run(static=False)
    got = run(static=True)
    for g, w, name, tol in zip(got, want, ("y", "running_mean", "running_var", "saved stats"),
                               (2e-2, 1e-4, 1e-4, 1e-3)):
        err = (g - w).abs()
        assert bool((err <= tol + tol * w.abs()).all()), f"{name}: max err {err.max().item():.4g}"
    # and against plain fp32 PyTorch, not only against the sibling entry point
    xf = x.float()
    ref = torch.nn.functional.batch_norm(xf, None, None, gamma, beta, True, 0.0, 1e-5)
    if res is not None:
        ref = ref + res.float()
    if relu:
        ref = torch.relu(ref)
    err = (got[0] - ref).abs()
    assert bool((err <= 3e-2 + 2e-2 * ref.abs()).all()), f"vs fp32 oracle: max err {err.max().item():.4g}"
    assert torch.allclose(got[1], 0.1 * xf.mean(0), atol=2e-3)                       # momentum 0.1 from zeros
    assert torch.allclose(got[2], 0.9 + 0.1 * xf.var(0, unbiased=True), rtol=2e-2, atol=2e-3)


def test_static_scratch_matches_caller_scratch_on_two_calls_in_a_row():
    lib = _abi()
    # first call: wide channel count fills a large part of the shared array with partial sums
    _case(lib, N=4, C=512, HW=14, relu=True, with_res=True, seed=1)
    # second call right behind it, different geometry: must not see the leftovers of the first
    _case(lib, N=8, C=64, HW=28, relu=False, with_res=False, seed=2)
    # and the first geometry again (same slots, new data)
    _case(lib, N=2, C=512, HW=7, relu=True, with_res=False, seed=3)


def test_static_scratch_rejects_sizes_beyond_its_limit():
    lib = _abi()
    limit = lib.cb2_bn_static_scratch_max_channels()
    assert limit == 2048
    dev = torch.device("cuda:0")
    C = limit * 2
    x = torch.zeros(8, C, device=dev, dtype=torch.bfloat16)
    buf = torch.zeros(2 * C, device=dev)
    rc = lib.cb2_bn_forward_static(x.data_ptr(), None, x.data_ptr(), buf.data_ptr(), buf.data_ptr(),
                                   buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), None,
                                   8, C, 1e-5, 0.1, 0, 1, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.cb2_last_error().decode()
    assert "2048" in msg and "scratch" in msg


def test_static_entry_point_inference_mode_and_argument_checks():
    """training = 0 applies the running statistics (the scratch is not involved); training = 2 - the
    caller-reduced-sums mode of cb2_bn_forward - is refused because there is no caller scratch."""
    lib = _abi()
    dev = torch.device("cuda:0")
    torch.manual_seed(4)
    M, C = 300, 64
    x = torch.randn(M, C, device=dev).to(torch.bfloat16)
    gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
    rm, rv = torch.randn(C, device=dev) * 0.1, torch.rand(C, device=dev) + 0.5
    y = torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    args = (gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), None, None, None, M, C, 1e-5, 0.1, 0)
    rc = lib.cb2_bn_forward_static(x.data_ptr(), None, y.data_ptr(), *args, 0, st)
    assert rc == 0, lib.cb2_last_error().decode()
    torch.cuda.synchronize()
    ref = torch.nn.functional.batch_norm(x.float(), rm, rv, gamma, beta, False, 0.0, 1e-5)
    assert (y.float() - ref).abs().max().item() < 3e-2 * max(1.0, ref.abs().max().item())
    rc = lib.cb2_bn_forward_static(x.data_ptr(), None, y.data_ptr(), *args, 2, st)
    assert rc != 0 and "training" in lib.cb2_last_error().decode()
