"""grl_l2_norm_static (csrc/train_ops.cu): the gradient-norm reduction with its scratch in a file-scope device array, for
callers that cannot allocate.  Collected last on purpose.  Checked through the C ABI against
grl_l2_norm with a caller workspace -- same launch shape, same summation order, so bit-identical -- on two calls in a row
(the array keeps the previous call's partial sums), and against an fp64 sum."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


def _norms(lib, _lib, x, ws, out_ws, out_static, st):
    _lib.check(lib.grl_l2_norm(x.data_ptr(), x.numel(), out_ws.data_ptr(), ws.data_ptr(), st), "grl_l2_norm")
    _lib.check(lib.grl_l2_norm_static(x.data_ptr(), x.numel(), out_static.data_ptr(), st), "grl_l2_norm_static")
    torch.cuda.synchronize()
    return float(out_ws), float(out_static)


def test_static_scratch_norm_equals_workspace_norm_on_two_calls_in_a_row(dev):
    from graph_representation_learning_b200 import _lib
    lib = _lib.load()
    st = _lib.stream_ptr(dev)
    gen = torch.Generator().manual_seed(3)
    # every buffer exists before the calls under test: the entry point itself allocates nothing
    big = torch.randn(3_141_593, generator=gen).to(dev)             # the model's flat gradient bucket is 3.1 M floats
    small = (torch.randn(1000, generator=gen) * 7).to(dev)          # fewer blocks than the call before: stale partials behind
    ws = torch.empty(lib.grl_reduce_workspace_bytes(), dtype=torch.uint8, device=dev)
    out_ws = torch.zeros(1, device=dev)
    out_static = torch.zeros(1, device=dev)
    for x in (big, small, big):
        a, b = _norms(lib, _lib, x, ws, out_ws, out_static, st)
        assert a == b, (x.numel(), a, b)
        want = float(x.double().pow(2).sum().sqrt())
        assert abs(b - want) <= 1e-6 * want


def test_static_scratch_norm_rejects_sizes_beyond_its_limit(dev):
    from graph_representation_learning_b200 import _lib
    lib = _lib.load()
    x = torch.ones(8, device=dev)
    out = torch.zeros(1, device=dev)
    rc = lib.grl_l2_norm_static(x.data_ptr(), (1 << 28) + 1, out.data_ptr(), _lib.stream_ptr(dev))   # rejected before any launch
    assert rc == -3
    with pytest.raises(_lib.GrlError, match="exceeds"):
        _lib.check(rc, "grl_l2_norm_static")
    assert lib.grl_l2_norm_static(x.data_ptr(), 0, out.data_ptr(), _lib.stream_ptr(dev)) == -1
    torch.cuda.synchronize()
    assert float(out) == 0.0
