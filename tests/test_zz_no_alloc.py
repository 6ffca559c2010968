"""gsb_depth_edge_loss_static (include/gsplatloc_b200.h): the depth + edge loss with its block partials in a file-scope
device array instead of a caller-provided scratch buffer, for callers that cannot allocate.  Called through the C ABI and
checked against gsb_depth_edge_loss on the same inputs, on two calls in a row (the second smaller than the first, so the
shared array still holds the first call's partials beyond what the second uses).  Same kernels and the same reduction
order, so the tolerance is the tests' tightest (1e-6 relative); the file sorts last so that `pytest -x` never hides
another test behind it."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_BLOCKS = 32768          # GSB_STATIC_LOSS_MAX_BLOCKS


@pytest.fixture(scope="module")
def L(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gsplatloc_b200 import _lib
    return _lib


def _inputs(C, W, H, seed):
    g = torch.Generator().manual_seed(seed)
    d = torch.rand(C, H, W, generator=g) * 3 + 0.5
    d[torch.rand(C, H, W, generator=g) < 0.1] = 0.0              # holes: the trainer's mask
    gt = d[0] + torch.randn(H, W, generator=g) * 0.05
    return d.cuda().contiguous(), gt.cuda().contiguous()


def _blocks(C, W, H):
    return C * ((W + 31) // 32) * ((H + 7) // 8)


def _both(L, C, W, H, seed):
    lib = L.lib()
    d, gt = _inputs(C, W, H, seed)
    w_d, w_s = 0.8 / (H * W), 0.2 / (H * W)
    st = torch.cuda.current_stream().cuda_stream
    sums_a = torch.full((C, 2), -1.0, device="cuda")
    sums_b = torch.full((C, 2), -2.0, device="cuda")
    v_a, v_b = torch.zeros_like(d), torch.ones_like(d)
    ws = torch.empty(3 * _blocks(C, W, H), device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    L.check(lib.gsb_depth_edge_loss_static(C, W, H, d.data_ptr(), gt.data_ptr(), 0, 1, w_d, w_s, sums_b.data_ptr(),
                                           v_b.data_ptr(), st), "depth_edge_loss_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    L.check(lib.gsb_depth_edge_loss(C, W, H, d.data_ptr(), gt.data_ptr(), 0, 1, w_d, w_s, sums_a.data_ptr(),
                                    v_a.data_ptr(), ws.data_ptr(), st), "depth_edge_loss")
    torch.cuda.synchronize()
    return sums_a.cpu().numpy(), v_a.cpu().numpy(), sums_b.cpu().numpy(), v_b.cpu().numpy(), free0 - free1


def test_static_scratch_loss_matches_the_workspace_entry_point_twice_in_a_row(L):
    _both(L, 1, 64, 16, seed=1)                                   # module load / symbol lookup out of the way
    for k, (C, W, H) in enumerate([(3, 333, 201), (2, 160, 96)]):  # ragged blocks, then fewer blocks than before
        sums_a, v_a, sums_b, v_b, grew = _both(L, C, W, H, seed=10 + k)
        assert np.all(np.isfinite(sums_b)) and np.all(sums_b > 0)
        assert np.allclose(sums_b, sums_a, rtol=1e-6, atol=0), (sums_a, sums_b)
        assert np.allclose(v_b, v_a, rtol=1e-6, atol=1e-12)
        assert np.abs(v_a).max() > 0
        assert grew <= 0, f"device memory in use grew by {grew} bytes across the call"


def test_static_scratch_loss_rejects_sizes_beyond_the_array(L):
    lib = L.lib()
    C, W, H = 5, 1920, 1080                                       # 5 * 60 * 135 = 40500 blocks
    assert _blocks(C, W, H) > MAX_BLOCKS
    x = torch.zeros(16, device="cuda")                            # never touched: the call must fail before any launch
    rc = lib.gsb_depth_edge_loss_static(C, W, H, x.data_ptr(), x.data_ptr(), 0, 1, 1.0, 1.0, x.data_ptr(), x.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 2                                                # GSB_ERR_ARG
    msg = lib.gsb_last_error().decode()
    assert "GSB_STATIC_LOSS_MAX_BLOCKS" in msg and "40500" in msg
    C = 4
    assert _blocks(C, W, H) <= MAX_BLOCKS                         # the documented limit for 1920x1080
    torch.cuda.synchronize()
