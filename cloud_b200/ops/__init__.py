"""Hot ops of cloud_b200, backed by hand-written sm_100a kernels.

CUDA tensors always go through ``libcloud_b200_kernels.so`` (tcgen05/TMEM GEMM fed by TMA, fused
cross-entropy, fused optimizers, peer-memory collectives); there is no silent PyTorch fallback on a
GPU.  CPU tensors use a plain PyTorch implementation of the same maths so that the strategy /
Keras / tuner layers are testable (and usable: OneDeviceStrategy on the CPU preset) without a GPU.

Reference parity: these are the ops the reference makes TensorFlow execute on the accelerator
(SURVEY.md section 2.5, K1-K12); the reference itself contains no kernels.
"""
from ._lib import KernelsUnavailable, kernels, runtime, check  # noqa: F401
from .dense import gemm, dense, dense_reference  # noqa: F401
from .loss import softmax_cross_entropy, softmax_cross_entropy_reference  # noqa: F401
from .elementwise import dropout, dropout_plan, relu_bwd_bias_grad, normalize_u8_images, normalize_u8_images_s2d  # noqa: F401
from .optim import FlatOptimizerState, optim_flat, optim_tick, OPT_CODES, make_hyper  # noqa: F401

from .bn import batch_norm_nhwc, bn_available, fork as bn_fork  # noqa: F401,E402
from .pool import max_pool_nhwc  # noqa: F401,E402
from .conv import conv1x1, supported as conv1x1_supported, mode as conv1x1_mode  # noqa: F401,E402
from .conv import stem_conv, stem_conv_s2d, stem_supported, conv_kxk, kxk_supported, conv3x3_own, subsample  # noqa: F401,E402
from .convbn import conv1x1_bn, supported as convbn_supported, DeferredConv  # noqa: F401,E402


def kernel_launches() -> int:
    """How many cloud_b200 kernels this process has launched (bench.py reports it)."""
    import importlib
    total = 0
    for name in ("dense", "loss", "elementwise", "optim", "bn", "pool", "convbn"):   # `dense` the module, not the op
        total += importlib.import_module(f"{__name__}.{name}").launches()
    return total + importlib.import_module("cloud_b200.parallel.comm").launches()


def native_available() -> bool:
    """True when the kernel library can be loaded (says nothing about a GPU being present)."""
    try:
        kernels()
        return True
    except KernelsUnavailable:
        return False
