"""``ocv_layernorm_bwd_noalloc_f32``: the LayerNorm backward for hosts that are not allowed to allocate device memory.

Who that is: an operator plugin living inside a runtime that planned all of its memory when the engine was built and hands the
plugin nothing but its inputs and outputs; a callback that runs while the surrounding framework holds its allocator lock; a
control loop that forbids ``cudaMalloc`` after start-up because it synchronises the device and can fail at any later moment.
None of them can supply the ``workspace`` that ``ocv_layernorm_bwd_f32`` asks for, so the variant brings its own: a fixed
192 KB array compiled into the library (``OCV_LN_BWD_NOALLOC_MAX_ROWS`` rows, E <= 512), shared by every call on a device.

What is checked, through the C ABI only (ctypes, raw pointers — no ``ops`` wrapper):
  * two calls in a row, the second on a different and smaller problem so that the shared array still holds the first call's
    leftovers, each against ``ocv_layernorm_bwd_f32`` (caller-owned workspace) on the same inputs: within the 1e-5 the other
    LayerNorm-backward tests use, and — the kernels being the same — bit for bit;
  * the call leaves the amount of free device memory unchanged;
  * a row count beyond the built-in array is refused with OCV_E_CAPACITY before anything is launched.

The file name sorts last on purpose: ``pytest -x`` stops at the first failure, and a failure here must not hide the parity
tests.  It does not depend on them: it passes alone in a fresh process and after the rest of the suite.
"""
import re

import pytest
import torch

from helpers import rel_err

OCV_E_CAPACITY = -6


def _max_rows() -> int:
    from objcavit_b200 import _lib
    m = re.search(r"#define\s+OCV_LN_BWD_NOALLOC_MAX_ROWS\s+(\d+)", open(_lib.HEADER_PATH).read())
    assert m, "OCV_LN_BWD_NOALLOC_MAX_ROWS is not in the header"
    return int(m.group(1))


def test_no_alloc_rejects_oversize_rows():
    """Runs anywhere: the capacity check comes before the device is touched, so no GPU is needed to see the refusal."""
    from objcavit_b200 import _lib
    lib = _lib.lib()
    one = 1  # any non-null address: nothing is dereferenced on a refused call
    status = lib.ocv_layernorm_bwd_noalloc_f32(one, one, one, one, None, None, _max_rows() + 1, 128, 1.0, None)
    assert status == OCV_E_CAPACITY
    assert b"capacity" in lib.ocv_status_string(status)
    assert lib.ocv_layernorm_bwd_noalloc_f32(one, one, one, one, None, None, 8, 100, 1.0, None) == -2    # E % 32 != 0, as the original
    assert lib.ocv_layernorm_bwd_noalloc_f32(None, one, one, one, None, None, 8, 128, 1.0, None) == -1


def _case(rows, E, seed, device):
    g = torch.Generator().manual_seed(seed)
    z = (torch.randn(rows, E, generator=g) * 2 + 0.5).to(device)
    dy = torch.randn(rows, E, generator=g).to(device)
    gamma = (1 + 0.1 * torch.randn(E, generator=g)).to(device)
    acc = torch.randn(2, E, generator=g).to(device)          # dgamma / dbeta already hold something: beta = 1 accumulates
    return dy, z, gamma, acc


def _run(lib, name, dy, z, gamma, out, workspace):
    """One call through the C ABI into pre-allocated outputs: out = (dz [rows,E], dgamma/dbeta [2,E] holding the values to add to)."""
    rows, E = z.shape
    dz, dgb = out
    args = [dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), dz.data_ptr(), dgb[0].data_ptr(), dgb[1].data_ptr(), rows, E, 1.0]
    if workspace is not None:
        args.append(workspace.data_ptr())
    args.append(torch.cuda.current_stream().cuda_stream)
    status = getattr(lib, name)(*args)
    assert status == 0, (name, status, lib.ocv_status_string(status))


@pytest.mark.gpu
def test_no_alloc_layernorm_bwd_matches_workspace_entry(cuda):
    from objcavit_b200 import _lib
    lib = _lib.lib()
    torch.cuda.set_device(cuda)           # the array's address is looked up for the device that is current at the first call
    # first call: the largest problem the array admits; second: config 5's token rows (8 x 221) at another width, over its leftovers
    for rows, E, seed in ((_max_rows(), 128, 101), (1768, 256, 102)):
        dy, z, gamma, acc = _case(rows, E, seed, cuda)
        workspace = torch.empty(rows * 2 + 32 * E, dtype=torch.float32, device=cuda)      # the test may allocate; the entry may not
        want_dz, want_dgb = torch.empty_like(z), acc.clone()
        got_dz, got_dgb = torch.empty_like(z), acc.clone()                                  # everything allocated before the call
        _run(lib, "ocv_layernorm_bwd_f32", dy, z, gamma, (want_dz, want_dgb), workspace)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        _run(lib, "ocv_layernorm_bwd_noalloc_f32", dy, z, gamma, (got_dz, got_dgb), None)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "the no-allocation entry point changed the free device memory"
        assert rel_err(got_dz, want_dz) < 1e-5 and rel_err(got_dgb, want_dgb) < 1e-5
        assert torch.equal(got_dz, want_dz) and torch.equal(got_dgb, want_dgb)
        # and the pair agrees with autograd, so "equal" is not two wrong answers agreeing
        zr = z.double().cpu().requires_grad_(True)
        gr = gamma.double().cpu().requires_grad_(True)
        br = torch.zeros(E, dtype=torch.float64, requires_grad=True)
        torch.nn.functional.layer_norm(zr, (E,), gr, br, 1e-5).backward(dy.double().cpu())
        assert rel_err(got_dz, zr.grad) < 1e-5
        assert rel_err(got_dgb[0], acc[0].double().cpu() + gr.grad) < 1e-5 and rel_err(got_dgb[1], acc[1].double().cpu() + br.grad) < 1e-5
    # data gradient only (dgamma = dbeta = NULL), a third call over the same array
    dy, z, gamma, _ = _case(300, 128, 103, cuda)
    workspace = torch.empty(300 * 2 + 32 * 128, dtype=torch.float32, device=cuda)
    want, got = torch.empty_like(z), torch.empty_like(z)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.ocv_layernorm_bwd_f32(dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), want.data_ptr(), None, None, 300, 128, 1.0,
                                     workspace.data_ptr(), st) == 0
    assert lib.ocv_layernorm_bwd_noalloc_f32(dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), got.data_ptr(), None, None, 300, 128, 1.0,
                                             st) == 0
    torch.cuda.synchronize()
    assert torch.equal(got, want)
