"""GPU: the decode-step streaming projections (cram_linear_decode, tcgen05 swap-AB) against an fp64 evaluation of
the HF formulas and against the library GEMM + glue kernels they replace in the decode step."""
import pytest
import torch
import torch.nn.functional as F

from cram_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _mk(m, n, k, dtype, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = torch.randn(m, k, device=DEV, generator=g).to(dtype)
    w = (torch.randn(n, k, device=DEV, generator=g) * k ** -0.5).to(dtype)
    return x, w


def _close(got, want64, dtype, k):
    """|got - exact| within one rounding of the result plus fp32 accumulation noise."""
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    err = (got.double() - want64).abs()
    bound = eps * want64.abs() + 1e-6 * k ** 0.5 + 1e-3 * eps
    assert bool((err <= bound * 1.01).all()), f"max excess {(err - bound).max().item():.3e}"


SHAPES = [(64, 4608, 3584), (64, 3584, 18944), (37, 200, 136), (1, 128, 64), (64, 1000, 4096), (5, 384, 8)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("m,n,k", SHAPES)
def test_plain_with_bias(dtype, m, n, k):
    x, w = _mk(m, n, k, dtype, 1)
    b = torch.randn(n, device=DEV).to(dtype)
    got = ops.linear_decode(x, w, b)
    _close(got, x.double() @ w.double().T + b.double(), dtype, k)
    lib = F.linear(x, w, b)
    assert (got == lib).float().mean().item() > 0.9  # same math as the library GEMM up to accumulation order
    again = ops.linear_decode(x, w, b)
    assert torch.equal(got, again)  # split-K partials are summed in a fixed order
    _close(ops.linear_decode(x, w), x.double() @ w.double().T, dtype, k)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("m,n,k", SHAPES)
def test_residual_add_in_place(dtype, m, n, k):
    x, w = _mk(m, n, k, dtype, 2)
    res = torch.randn(m, n, device=DEV).to(dtype)
    got = res.clone()
    ops.linear_decode(x, w, mode=ops.LINEAR_RESIDUAL, out=got)
    exact = (x.double() @ w.double().T)
    o = exact.to(dtype)  # the projection rounds to the model dtype before the add (two roundings, as HF eager)
    want = res + o
    # where the fp32 accumulation rounds the projection differently from fp64, the sum moves by <= 1 ulp of o
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert bool(((got.double() - want.double()).abs() <= 2 * eps * (o.double().abs() + want.double().abs()) + 1e-4).all())
    assert (got == want).float().mean().item() > 0.9
    lib = res + F.linear(x, w)
    assert (got == lib).float().mean().item() > 0.9


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("m,inter,k", [(64, 18944, 3584), (64, 14336, 4096), (19, 104, 72), (1, 64, 64), (64, 512, 256)])
def test_swiglu(dtype, m, inter, k):
    x, w = _mk(m, 2 * inter, k, dtype, 3)
    got = ops.linear_decode(x, w, mode=ops.LINEAR_SWIGLU)
    gu = (x.double() @ w.double().T).to(dtype)
    want = (F.silu(gu[:, :inter]) * gu[:, inter:])
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert bool(((got.double() - want.double()).abs() <= 8 * eps * want.double().abs() + 1e-4).all())
    lib = ops.silu_mul(F.linear(x, w))
    assert (got == lib).float().mean().item() > 0.9
    assert got.shape == (m, inter)


def test_many_launches_give_the_same_bits():
    x, w = _mk(64, 3584, 3584, torch.bfloat16, 4)
    first = ops.linear_decode(x, w)
    for _ in range(20):
        assert torch.equal(ops.linear_decode(x, w), first)


def test_rejects_more_than_64_rows():
    x, w = _mk(65, 128, 64, torch.bfloat16, 5)
    with pytest.raises(RuntimeError):
        ops.linear_decode(x, w)
