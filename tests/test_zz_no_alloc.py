"""The allocation-free pool-branch entry point against the allocating one, through the C ABI.

Some users of the library may not allocate GPU memory at the moment they call it: a Flink task thread that shares
its GPU with other subtasks (a device allocation or free can stall everybody's streams), or a job whose HBM was
fully handed out when the operators opened.  ftf_pool_branch_bf16_static_dev serves them from a fixed 8 MiB scratch
array inside the library.  It must give exactly what ftf_pool_branch_bf16_dev gives, also on a second call that
finds the first call's leftovers in the scratch, and must refuse work that does not fit.

This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(lib, name, x, w, b, out):
    import torch
    n, h, wd, cin = x.shape
    fn = getattr(lib, name)
    return fn(x.data_ptr(), n, h, wd, cin, cin, w.data_ptr(), b.data_ptr(), w.shape[0], out.data_ptr(), out.shape[-1], 0,
              C.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_static_scratch_pool_branch_matches_allocating_entry_point(ftf):
    import torch
    lib = ftf._abi.load_library()
    g = torch.Generator(device="cuda").manual_seed(99)
    # two calls in a row; the second is smaller, so it runs over scratch that still holds the first call's values
    for n, h, wd, cin, cout in [(3, 17, 17, 768, 192), (2, 8, 8, 1280, 192)]:
        x = torch.randn((n, h, wd, cin), device="cuda", generator=g).to(torch.bfloat16)
        w = (torch.randn((cout, cin), device="cuda", generator=g) * 0.05).to(torch.bfloat16)
        b = torch.rand(cout, device="cuda", generator=g) * 0.2 - 0.1
        want = torch.full((n, h, wd, cout), -1.0, dtype=torch.bfloat16, device="cuda")
        got = torch.full((n, h, wd, cout), -2.0, dtype=torch.bfloat16, device="cuda")
        assert _call(lib, "ftf_pool_branch_bf16_dev", x, w, b, want) == 0, lib.ftf_last_error(None)
        assert _call(lib, "ftf_pool_branch_bf16_static_dev", x, w, b, got) == 0, lib.ftf_last_error(None)
        torch.cuda.synchronize()
        # same kernels, same arithmetic, same order: the tolerance is zero
        assert torch.equal(got, want)
        assert float(want.float().abs().max()) > 0.0


def test_static_scratch_refuses_what_does_not_fit(ftf):
    import torch
    lib = ftf._abi.load_library()
    n, h, wd, cin, cout = 64, 35, 35, 64, 64                      # 64*35*35*64*4 B = 20 MB > 8 MiB
    x = torch.zeros((n, h, wd, cin), dtype=torch.bfloat16, device="cuda")
    w = torch.zeros((cout, cin), dtype=torch.bfloat16, device="cuda")
    b = torch.zeros(cout, device="cuda")
    out = torch.full((n, h, wd, cout), 5.0, dtype=torch.bfloat16, device="cuda")
    rc = _call(lib, "ftf_pool_branch_bf16_static_dev", x, w, b, out)
    assert rc == ftf._abi.RESOURCE_EXHAUSTED
    assert b"exceeds the built-in scratch" in lib.ftf_last_error(None)
    torch.cuda.synchronize()
    assert torch.all(out == 5.0)                                   # nothing was launched
