"""GPU test of `ai4e_head_noalloc`, the head for callers that may not allocate device memory
(their framework owns all of the GPU's memory, or they cannot afford cudaMalloc's device-wide
synchronisation).  It must give what `ai4e_head` gives - same kernels, only the scratch differs -
on two calls in a row that share the library's fixed scratch array, and refuse a batch that
does not fit that array.  Collected last on purpose (file name), so a failure here cannot hide
the result of any other test."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(b, hw, c, classes, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.relu(torch.randn((b, hw, hw, c), generator=g)).to(torch.bfloat16).cuda()
    w = (torch.randn((classes, c), generator=g) * 0.02).to(torch.bfloat16).cuda()
    bias = (torch.randn((classes,), generator=g) * 0.1).cuda()
    return x, w, bias


def _same(a, b):
    la, ia, pa = a
    lb, ib, pb = b
    # the tolerance the head's own parity test uses (tests/test_gpu_kernels.py::test_head_vs_oracle)
    assert torch.allclose(la, lb, atol=2e-4, rtol=1e-4), (la - lb).abs().max()
    assert torch.equal(ia, ib)
    assert torch.allclose(pa, pb, atol=1e-6, rtol=1e-4)


def test_head_noalloc_matches_head_twice_in_a_row():
    from ai4e_b200 import ops
    # first call fills the shared scratch with a 48-image batch ...
    x1, w1, b1 = _case(48, 7, 2048, 1000, seed=11)
    want1 = ops.head(x1, w1, b1)
    got1 = ops.head_noalloc(x1, w1, b1)
    torch.cuda.synchronize()
    _same(got1, want1)
    # on its own feet too: the logits against plain fp32 torch (mean over the 7x7 positions, then FC)
    ref = x1.float().mean(dim=(1, 2)) @ w1.float().t() + b1
    assert torch.allclose(got1[0], ref, atol=2e-4, rtol=1e-4), (got1[0] - ref).abs().max()
    assert torch.equal(got1[1][:, 0].long(), got1[0].argmax(dim=1))
    # ... the second, smaller and different, must not see anything of the first
    x2, w2, b2 = _case(5, 8, 512, 1000, seed=12)
    want2 = ops.head(x2, w2, b2)
    got2 = ops.head_noalloc(x2, w2, b2)
    torch.cuda.synchronize()
    _same(got2, want2)
    # and the first again, bit for bit what it gave before
    again = ops.head_noalloc(x1, w1, b1)
    torch.cuda.synchronize()
    assert torch.equal(again[0], got1[0]) and torch.equal(again[1], got1[1])


def test_head_noalloc_refuses_what_does_not_fit():
    from ai4e_b200 import ops, Ai4eError, _lib as L
    b = L.HEAD_NOALLOC_MAX_FLOATS // 2048 + 1
    x, w, bias = _case(b, 1, 2048, 16, seed=13)
    with pytest.raises(Ai4eError) as ei:
        ops.head_noalloc(x, w, bias)
    assert ei.value.code == L.ERR_INVALID and "scratch" in str(ei.value)
    # exactly at the limit is fine
    x, w, bias = _case(b - 1, 1, 2048, 16, seed=14)
    _same(ops.head_noalloc(x, w, bias), ops.head(x, w, bias))
