"""lgk_colsum_static (include/lgk.h): the bias-gradient column sum with its scratch in a device array inside the library,
for callers that cannot allocate.  No counterpart in the reference.
Called through the C ABI and checked against the existing entry point lgk_linear_bwd_w(dw = NULL) on two calls in a row
(the second smaller than the first, so stale scratch contents would show), in fp32 and bf16, plain and accumulating; the
size limit is rejected with a clear error; torch's allocator sees no allocation during the calls.

This file is named to be collected LAST."""
import pytest
import torch

from tests.util import TOL_BF16, TOL_F32, assert_close

pytestmark = pytest.mark.gpu

MAX_M = 1024  # LGK_COLSUM_STATIC_MAX_M


def _existing(dy, db_out=None):
    """the existing entry point: lgk_linear_bwd_w with dw = NULL and a caller-provided workspace"""
    from lesion_gnn_b200 import ops
    return ops.linear_bwd_w(dy, None, want_dw=False, want_db=True, db_out=db_out)[1]


def _static(dy, db, accumulate):
    from lesion_gnn_b200 import _lib
    L = _lib.lib()
    n, m = dy.shape
    dt = _lib.F32 if dy.dtype == torch.float32 else _lib.BF16
    L.call("lgk_colsum_static", dy.data_ptr(), db.data_ptr(), n, m, int(accumulate), dt, torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_column_sum_equals_the_workspace_one(dev, dtype):
    gen = torch.Generator(device=dev).manual_seed(5)
    shapes = [(40000, MAX_M), (777, 64), (11447, 260), (1, 5), (65, 3)]  # big first: later calls find its leftovers
    tol = TOL_F32 if dtype == torch.float32 else TOL_BF16
    for n, m in shapes:
        dy = torch.randn(n, m, device=dev, generator=gen).to(dtype).contiguous()
        want = _existing(dy)
        exact = dy.double().sum(0)
        for call in range(2):  # two calls in a row on the same scratch
            db = torch.full((m,), float("nan"), device=dev)
            before = torch.cuda.memory_allocated(dev)
            _static(dy, db, accumulate=False)
            assert torch.cuda.memory_allocated(dev) == before
            assert_close(db, want, tol, f"n={n} m={m} call {call} vs lgk_linear_bwd_w")
            assert torch.equal(db, want), "same kernels, same order: bit-identical"
            assert_close(db, exact, TOL_F32, f"n={n} m={m} vs the fp64 sum of the same (rounded) inputs: fp32 accumulation")
        # accumulating form: db += colsum
        base = torch.randn(m, device=dev, generator=gen)
        acc_want = base.clone()
        _existing(dy, db_out=acc_want)
        acc = base.clone()
        _static(dy, acc, accumulate=True)
        assert torch.equal(acc, acc_want)


def test_static_scratch_empty_input_and_limit(dev):
    from lesion_gnn_b200 import _lib
    L = _lib.lib()
    s = torch.cuda.current_stream().cuda_stream
    db = torch.full((8,), 3.0, device=dev)
    L.call("lgk_colsum_static", None, db.data_ptr(), 0, 8, 0, _lib.F32, s)  # no rows: the empty sum
    assert float(db.abs().max()) == 0.0
    dy = torch.zeros(4, MAX_M + 4, device=dev)
    out = torch.zeros(MAX_M + 4, device=dev)
    with pytest.raises(RuntimeError, match="LGK_COLSUM_STATIC_MAX_M"):
        L.call("lgk_colsum_static", dy.data_ptr(), out.data_ptr(), 4, MAX_M + 4, 0, _lib.F32, s)
    with pytest.raises(RuntimeError):
        L.call("lgk_colsum_static", dy.data_ptr(), out.data_ptr(), 4, 8, 0, 7, s)  # unknown dtype
