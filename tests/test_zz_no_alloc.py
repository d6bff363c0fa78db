"""dlrmx_relu_bwd_bias_grad_noalloc (the entry point that allocates nothing: its scratch is a static __device__ array
inside libdlrmx.so) against dlrmx_relu_bwd_bias_grad with a caller-provided workspace, both called through the C ABI.
Same kernel, same grid: the outputs must be bit-identical.  Two calls in a row (the second on other data and another
shape) check that the shared scratch, which keeps the previous call's contents, is cleared where it must be.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _existing(lib, dy, y, relu):
    """The existing entry point with the wrapper's default chunking and caller-allocated partials / counters."""
    from dlrm_yx_b200 import _lib
    B, N = dy.shape
    col_blocks = (N + 255) // 256
    chunks = max(1, min((B + 31) // 32, (148 * 2 + col_blocks - 1) // col_blocks))
    dz = torch.empty_like(dy)
    part = torch.empty((chunks, N), dtype=torch.float32, device=dy.device)
    counters = torch.zeros(64, dtype=torch.int32, device=dy.device)
    db = torch.full((N,), float("nan"), device=dy.device)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.dlrmx_relu_bwd_bias_grad(dy.data_ptr(), y.data_ptr(), dz.data_ptr(), part.data_ptr(), db.data_ptr(),
                                      counters.data_ptr(), B, N, chunks, int(relu), st)
    _lib.check(rc, "dlrmx_relu_bwd_bias_grad")
    return dz, db


def _noalloc(lib, dy, y, relu):
    from dlrm_yx_b200 import _lib
    B, N = dy.shape
    dz = torch.empty_like(dy)                       # outputs are the caller's
    db = torch.full((N,), float("nan"), device=dy.device)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.dlrmx_relu_bwd_bias_grad_noalloc(dy.data_ptr(), y.data_ptr(), dz.data_ptr(), db.data_ptr(), B, N, int(relu),
                                              torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "dlrmx_relu_bwd_bias_grad_noalloc")
    assert torch.cuda.memory_allocated() == before  # nothing came from torch's allocator either
    return dz, db


def test_noalloc_entry_point_matches_existing_entry_point_twice_in_a_row():
    from dlrm_yx_b200 import _lib
    lib = _lib.load()
    torch.manual_seed(21)
    for B, N in [(32768, 1024), (1000, 264)]:       # second call: other data, other shape, stale scratch from the first
        y = torch.relu(torch.randn(B, N, device="cuda")).to(torch.bfloat16)
        dy = torch.randn(B, N, device="cuda").to(torch.bfloat16)
        dz_ref, db_ref = _existing(lib, dy, y, relu=True)
        dz, db = _noalloc(lib, dy, y, relu=True)
        torch.cuda.synchronize()
        assert torch.equal(dz, dz_ref)
        assert torch.equal(db, db_ref)               # same kernel, same grid, same summation order: bit-identical
        want = dy.float() * (y.float() > 0)
        torch.testing.assert_close(db, want.sum(0), rtol=1e-4, atol=1e-3)
    # without the ReLU: column sums of dy, dz / y untouched
    dz0 = torch.zeros_like(dy)
    db0 = torch.empty(N, device="cuda")
    rc = lib.dlrmx_relu_bwd_bias_grad_noalloc(dy.data_ptr(), None, None, db0.data_ptr(), B, N, 0,
                                              torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "dlrmx_relu_bwd_bias_grad_noalloc")
    _, db_ref0 = _existing(lib, dy, y, relu=False)
    torch.cuda.synchronize()
    assert torch.equal(db0, db_ref0)
    assert float(dz0.abs().max()) == 0.0


def test_noalloc_entry_point_rejects_sizes_beyond_its_static_scratch():
    from dlrm_yx_b200 import _lib
    lib = _lib.load()
    B, N = 64, 2048 + 8
    t = torch.zeros(B, N, dtype=torch.bfloat16, device="cuda")
    db = torch.zeros(N, device="cuda")
    rc = lib.dlrmx_relu_bwd_bias_grad_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), db.data_ptr(), B, N, 1,
                                              torch.cuda.current_stream().cuda_stream)
    assert rc == -1                                  # DLRMX_E_ARG
    assert b"exceeds the static scratch" in lib.dlrmx_last_error()
    # the limit itself is accepted
    t2 = torch.zeros(B, 2048, dtype=torch.bfloat16, device="cuda")
    rc = lib.dlrmx_relu_bwd_bias_grad_noalloc(t2.data_ptr(), t2.data_ptr(), t2.data_ptr(), db.data_ptr(), B, 2048, 1,
                                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
