"""The allocation-free MSE entry point, called through the C ABI.

``b200dt_mse_loss_static`` exists for callers that are not allowed to allocate device memory
(their memory was partitioned up front, or another runtime owns the heap): it keeps its reduction
scratch in a fixed array inside the library instead of taking a scratch buffer.  This test drives
it with ctypes exactly as a C program would, twice in a row on the same stream (the second call
meets the array as the first one left it), and compares loss and gradient with the ordinary
``b200dt_mse_loss`` entry point, which is itself checked against PyTorch in test_kernels_gpu.py.
The file name sorts last on purpose: a run that stops at the first failure has already executed
every other test by the time it gets here.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1048576          # B200DT_MSE_STATIC_MAX_ELEMS in include/b200dt.h


def _abi():
    from pytorch_distributed_training_tutorials_b200.build import lib_path
    lib = C.CDLL(str(lib_path()))
    vp, i64, f32, i32 = C.c_void_p, C.c_longlong, C.c_float, C.c_int
    lib.b200dt_mse_loss.argtypes = [vp, vp, vp, vp, i64, f32, i32, vp, vp]
    lib.b200dt_mse_loss.restype = i32
    lib.b200dt_mse_loss_static.argtypes = [vp, vp, vp, vp, i64, f32, i32, vp]
    lib.b200dt_mse_loss_static.restype = i32
    lib.b200dt_loss_scratch_bytes.restype = i64
    return lib


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [1, 4097, 120 * 1000, MAX_ELEMS])
def test_static_scratch_mse_matches_regular_entry_point(dtype, n):
    lib = _abi()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    is_bf16 = int(dtype == torch.bfloat16)
    scratch = torch.zeros(int(lib.b200dt_loss_scratch_bytes()), dtype=torch.uint8, device=dev)
    gen = torch.Generator(device="cpu").manual_seed(n)
    for call in range(2):                       # two calls in a row share the static array
        pred = torch.randn(n, generator=gen).to(dev, dtype)
        target = torch.randn(n, generator=gen).to(dev, dtype)
        grad_ref = torch.empty_like(pred); loss_ref = torch.zeros((), device=dev)
        grad_new = torch.empty_like(pred); loss_new = torch.full((), -1.0, device=dev)
        rc = lib.b200dt_mse_loss(pred.data_ptr(), target.data_ptr(), grad_ref.data_ptr(), loss_ref.data_ptr(),
                                 n, 0.5, is_bf16, scratch.data_ptr(), stream)
        assert rc == 0
        rc = lib.b200dt_mse_loss_static(pred.data_ptr(), target.data_ptr(), grad_new.data_ptr(),
                                        loss_new.data_ptr(), n, 0.5, is_bf16, stream)
        assert rc == 0, f"call {call}: rc={rc}"
        torch.cuda.synchronize(dev)
        # same per-element arithmetic: the gradients agree exactly; the scalar differs only by
        # the order in which partial sums are added
        assert torch.equal(grad_new, grad_ref), call
        ref = loss_ref.item()
        assert abs(loss_new.item() - ref) <= 1e-5 * max(1.0, abs(ref)), (call, loss_new.item(), ref)


def test_static_scratch_mse_rejects_oversize_and_null():
    lib = _abi()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    x = torch.zeros(8, device=dev); loss = torch.full((), 7.0, device=dev)
    rc = lib.b200dt_mse_loss_static(x.data_ptr(), x.data_ptr(), None, loss.data_ptr(), MAX_ELEMS + 1, 1.0, 0, stream)
    assert rc == -9                             # beyond the static scratch: refused, nothing launched
    rc = lib.b200dt_mse_loss_static(None, x.data_ptr(), None, loss.data_ptr(), 8, 1.0, 0, stream)
    assert rc == -3
    torch.cuda.synchronize(dev)
    assert loss.item() == 7.0
    # gradient output is optional
    rc = lib.b200dt_mse_loss_static(x.data_ptr(), x.data_ptr(), None, loss.data_ptr(), 8, 1.0, 0, stream)
    assert rc == 0
    torch.cuda.synchronize(dev)
    assert loss.item() == 0.0
