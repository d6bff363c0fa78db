"""tcgen05 tower GEMMs (csrc/gemm_tc.cu) through the C ABI against fp32 torch matmuls of the SAME bf16-rounded operands.
bf16 outputs: one bf16 rounding of an fp32-accumulated value -> |got - ref| <= 2^-8 |ref| + 1e-5 * (sum_k |a||b|) per
element (accumulation order differs from torch's); fp32 weight gradients: 1e-4 of the product-magnitude sum.  Gate bits
must equal (stored bf16 output > 0) exactly.  Ragged shapes: M not a multiple of the 256-row pair tile, K and N not
multiples of the 64 / 128 / 256 tile extents, single rows, widths of 8."""
import pytest
import torch

pytestmark = pytest.mark.gpu

BF = torch.bfloat16


@pytest.fixture(autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.backends.cuda.matmul.allow_tf32 = False


def _operands(M, K, N, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.randn(M, K, device="cuda", generator=g) * 0.5).to(BF)
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).to(BF)
    b = torch.randn(N, device="cuda", generator=g)
    return x, w, b


def _close_bf16(got, ref, mag):
    err = (got.float() - ref).abs()
    bound = 2.0 ** -8 * ref.abs() + 1e-5 * mag + 1e-7
    assert bool((err <= bound).all()), float((err - bound).max())


def _unpack_bits(bits, N):
    sh = torch.arange(32, device=bits.device, dtype=torch.int32)
    return ((bits[:, :, None] >> sh) & 1).reshape(bits.shape[0], -1)[:, :N].bool()


SHAPES = [(1, 8, 8), (257, 16, 512), (4099, 480, 1024), (300, 104, 64), (1000, 512, 256), (513, 1024, 136), (129, 72, 40)]


@pytest.mark.parametrize("M,K,N", SHAPES)
def test_forward_bias_relu_and_gate_bits(M, K, N):
    from dlrm_yx_b200 import ops
    x, w, b = _operands(M, K, N, 1)
    acc = x.float() @ w.float().t()
    mag = x.float().abs() @ w.float().abs().t()
    bits = torch.full((M, (N + 31) // 32), -1, dtype=torch.int32, device="cuda")
    y = ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS_RELU, bias=b, relu_bits_out=bits)
    _close_bf16(y, torch.relu(acc + b), mag + b.abs())
    assert torch.equal(_unpack_bits(bits, N), y.float() > 0)          # the gates are those of the STORED bf16 output
    _close_bf16(ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS, bias=b), acc + b, mag + b.abs())
    _close_bf16(ops.gemm_bf16(x, w), acc, mag)
    # a strided output (a column window of a wider matrix) and a strided A
    wide = torch.zeros(M, N + 16, dtype=BF, device="cuda")
    xa = torch.zeros(M, K + 8, dtype=BF, device="cuda")
    xa[:, :K] = x
    ops.gemm_bf16(xa[:, :K], w, out=wide[:, 8:8 + N])
    _close_bf16(wide[:, 8:8 + N], acc, mag)
    assert float(wide[:, :8].abs().max()) == 0.0 and float(wide[:, 8 + N:].abs().max()) == 0.0


@pytest.mark.parametrize("M,K,N", SHAPES)
def test_dgrad_with_gate_bits(M, K, N):
    """dx [M, K] = (dz [M, N] . W [N, K]) gated by the previous layer's output: W is the MN-major B operand."""
    from dlrm_yx_b200 import ops
    _, w, _ = _operands(M, K, N, 2)
    g = torch.Generator(device="cuda").manual_seed(3)
    dz = (torch.randn(M, N, device="cuda", generator=g) * 0.1).to(BF)
    yprev = torch.relu(torch.randn(M, K, device="cuda", generator=g)).to(BF)
    acc = dz.float() @ w.float()
    mag = dz.float().abs() @ w.float().abs()
    gate = yprev.float() > 0
    # bits as the forward writes them
    sh = torch.arange(32, device="cuda", dtype=torch.int64)
    padded = torch.zeros(M, (K + 31) // 32 * 32, dtype=torch.int64, device="cuda")
    padded[:, :K] = gate
    words = (padded.view(M, -1, 32) << sh).sum(-1)
    bits = torch.where(words >= 2 ** 31, words - 2 ** 32, words).to(torch.int32)
    dx = ops.gemm_bf16(dz, w, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=bits)
    _close_bf16(dx, acc * gate, mag)
    assert bool((dx.float()[~gate] == 0).all())
    _close_bf16(ops.gemm_bf16(dz, w, b_is_kn=True), acc, mag)


@pytest.mark.parametrize("M,layers", [(257, [(16, 512), (512, 256), (256, 128)]),
                                      (4099, [(480, 1024), (1024, 1024), (1024, 512), (512, 256)]),
                                      (1000, [(16, 512), (256, 64), (64, 16), (368, 512)]),
                                      (64, [(8, 8)]), (70000, [(72, 40), (104, 1024)])])
def test_wgrad_all_layers_one_launch(M, layers):
    """dW_l = dZ_l^t X_l and db_l = column sums of dZ_l for every layer in one split-K launch (+ one reduce launch);
    deterministic (two calls, bit-equal); a layer with db = None gets no bias gradient."""
    from dlrm_yx_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(4)
    dzs = [(torch.randn(M, N, device="cuda", generator=g) * 0.1).to(BF) for K, N in layers]
    xs = [(torch.randn(M, K, device="cuda", generator=g) * 0.5).to(BF) for K, N in layers]
    dws = [torch.full((N, K), float("nan"), device="cuda") for K, N in layers]
    dbs = [torch.full((N,), float("nan"), device="cuda") for K, N in layers]
    dbs[-1] = None
    ws = ops.gemm_wgrad_workspace([N for K, N in layers], [K for K, N in layers], M, torch.device("cuda"))
    ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, ws)
    first = [d.clone() for d in dws]
    for dz, x, dw, db in zip(dzs, xs, dws, dbs):
        ref = dz.float().t() @ x.float()
        mag = dz.float().abs().t() @ x.float().abs()
        assert bool(((dw - ref).abs() <= 1e-4 * mag + 1e-6).all()), float((dw - ref).abs().max())
        if db is not None:
            torch.testing.assert_close(db, dz.float().sum(0), rtol=1e-4, atol=1e-4 * float(dz.float().abs().sum(0).max()))
    ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, ws)
    assert all(torch.equal(a, b) for a, b in zip(first, dws))
