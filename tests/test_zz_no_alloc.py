"""yat_colsum_bf16_static: the column sum (bias gradient of a Linear) with its split partials in a device array built into the
library, for callers that cannot allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
Called through the C-ABI; must equal yat_colsum_bf16 (same kernels, same fixed
summation order -> bit-identical, and within the 4e-3 of tests/test_block_ops_gpu.py of the fp64 sum) on two calls in a
row that share the scratch array.  This file is collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-300)).item()


def _existing(L, a, st):
    from yat_b200 import _lib
    rows, C = a.shape
    out = torch.empty((C,), dtype=torch.bfloat16, device=a.device)
    ws = torch.empty(L.yat_colsum_workspace_bytes(rows, C), dtype=torch.uint8, device=a.device)
    _lib.check(L.yat_colsum_bf16(a.data_ptr(), rows, C, out.data_ptr(), ws.data_ptr(), st), "yat_colsum_bf16")
    return out


def test_static_scratch_colsum_matches_the_workspace_entry_point(cuda):
    from yat_b200 import _lib
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    assert L.yat_colsum_static_max_bytes() == 2 << 20
    g = torch.Generator().manual_seed(5)
    # two calls in a row on the shared array: a large shape (132 splits x 2240 columns of partials), then a smaller one
    # whose partials overwrite only the head of what the first call left behind
    shapes = [(16384, 2240), (777, 72), (100, 11200), (1, 8)]
    inputs = [torch.randn(s, generator=g).to(torch.bfloat16).to(cuda) for s in shapes]
    outs = [torch.full((s[1],), float("nan"), dtype=torch.bfloat16, device=cuda) for s in shapes]
    L.yat_colsum_bf16_static(inputs[0].data_ptr(), 8, 8, outs[3].data_ptr(), st)      # resolves the device address (set-up)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    alloc0 = torch.cuda.memory_allocated()
    for a, o in zip(inputs, outs):                                                     # back to back, no sync in between
        _lib.check(L.yat_colsum_bf16_static(a.data_ptr(), a.shape[0], a.shape[1], o.data_ptr(), st), "yat_colsum_bf16_static")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0 and torch.cuda.memory_allocated() == alloc0   # nothing was allocated
    for a, o in zip(inputs, outs):
        ref = _existing(L, a, st)
        torch.cuda.synchronize()
        exact = a.double().sum(0)
        assert torch.equal(o, ref)
        assert torch.equal(o, exact.to(torch.bfloat16)) or _rel(o, exact) < 4e-3
    # and again, in the other order, into fresh outputs
    for a in reversed(inputs):
        o = torch.full((a.shape[1],), float("nan"), dtype=torch.bfloat16, device=cuda)
        _lib.check(L.yat_colsum_bf16_static(a.data_ptr(), a.shape[0], a.shape[1], o.data_ptr(), st), "yat_colsum_bf16_static")
        assert torch.equal(o, _existing(L, a, st))


def test_static_scratch_rejects_what_does_not_fit(cuda):
    from yat_b200 import _lib
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    a = torch.zeros((8, 8), dtype=torch.bfloat16, device=cuda)
    o = torch.zeros((8,), dtype=torch.bfloat16, device=cuda)
    C = (2 << 20) // 4 + 8                                   # one split of this width already exceeds the array
    assert L.yat_colsum_workspace_bytes(1, C) > L.yat_colsum_static_max_bytes()
    rc = L.yat_colsum_bf16_static(a.data_ptr(), 1, C, o.data_ptr(), st)          # rejected before anything is read
    assert rc != 0
    with pytest.raises(Exception, match="scratch"):
        _lib.check(rc, "yat_colsum_bf16_static")
    with pytest.raises(Exception):
        _lib.check(L.yat_colsum_bf16_static(a.data_ptr(), 8, 12, o.data_ptr(), st), "yat_colsum_bf16_static")   # C % 8
    with pytest.raises(Exception):
        _lib.check(L.yat_colsum_bf16_static(None, 8, 8, o.data_ptr(), st), "yat_colsum_bf16_static")
