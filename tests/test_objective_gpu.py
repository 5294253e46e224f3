"""Fused NerfWLoss kernels (csrc/objective.cu) against the torch-op restatement of losses.py:35-50 and the oracle."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(n, S, seed, dev):
    g = torch.Generator().manual_seed(seed)
    res = {"rgb_coarse": torch.rand(n, 3, generator=g), "rgb_fine": torch.rand(n, 3, generator=g),
           "beta": torch.rand(n, generator=g) * 0.5 + 0.05, "transient_sigmas": torch.rand(n, S, generator=g) * 3}
    tgt = torch.rand(n, 3, generator=g)
    return {k: v.to(dev) for k, v in res.items()}, tgt.to(dev)


@pytest.mark.parametrize("n,S", [(1, 1), (7, 5), (1024, 128), (4099, 33)])
@pytest.mark.parametrize("variant", ["nerfw", "no_transient", "coarse_only"])
def test_nerfw_loss_matches_torch_ops(n, S, variant):
    from nerf_fl_b200.objective import NerfWLoss
    from tests.host_models import nerfw_terms
    dev = torch.device("cuda")
    res, tgt = _inputs(n, S, 11 * n + S, dev)
    if variant != "nerfw":
        res.pop("beta"); res.pop("transient_sigmas")
    if variant == "coarse_only":
        res.pop("rgb_fine")
    a = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    b = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    coef, lam = 1.5, 0.02
    got = NerfWLoss(coef=coef, lambda_u=lam)(a, tgt)
    want = {k: coef * v for k, v in nerfw_terms(b, tgt, lam).items()}
    assert set(got) == set(want)
    for k in want:
        torch.testing.assert_close(got[k], want[k], rtol=2e-5, atol=1e-6)
    torch.testing.assert_close(got.total, sum(want.values()), rtol=2e-5, atol=1e-6)
    # cotangents: through the per-term outputs with unequal weights, and through `total`
    w = {"c_l": 1.0, "f_l": 0.7, "b_l": -0.3, "s_l": 2.0}
    sum(w[k] * got[k] for k in got).backward()
    sum(w[k] * want[k] for k in want).backward()
    for k in a:
        torch.testing.assert_close(a[k].grad, b[k].grad, rtol=2e-5, atol=1e-8)
    a2 = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    b2 = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    NerfWLoss(coef=coef, lambda_u=lam)(a2, tgt).total.backward()
    sum(coef * v for v in nerfw_terms(b2, tgt, lam).values()).backward()
    for k in a2:
        torch.testing.assert_close(a2[k].grad, b2[k].grad, rtol=2e-5, atol=1e-8)


def test_nerfw_loss_matches_oracle():
    from nerf_fl_b200.objective import NerfWLoss
    from oracle import nerfw_oracle as O
    dev = torch.device("cuda")
    res, tgt = _inputs(513, 17, 5, dev)
    got = NerfWLoss(coef=1, lambda_u=0.01)(res, tgt)
    want = O.nerfw_loss({k: v.cpu() for k, v in res.items()}, tgt.cpu(), coef=1, lambda_u=0.01)
    for k in want:
        torch.testing.assert_close(got[k].cpu(), want[k], rtol=2e-5, atol=1e-6)


def test_nerfw_loss_is_deterministic_and_reusable():
    from nerf_fl_b200.objective import NerfWLoss
    res, tgt = _inputs(2048, 64, 3, torch.device("cuda"))
    f = NerfWLoss()
    first = f(res, tgt).total.item()
    for _ in range(5):
        assert f(res, tgt).total.item() == first


@pytest.mark.parametrize("with_fine", [True, False])
def test_color_loss_matches_torch_ops(with_fine):
    """losses.py:4-16 through the same kernel pair (no beta: total = coef' * 0.5 * (mse_c + mse_f), coef' = 2 coef)."""
    from nerf_fl_b200.objective import ColorLoss
    from tests.host_models import color_loss
    res, tgt = _inputs(777, 4, 9, torch.device("cuda"))
    res = {k: v for k, v in res.items() if k == "rgb_coarse" or (with_fine and k == "rgb_fine")}
    a = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    b = {k: v.clone().requires_grad_(True) for k, v in res.items()}
    got, want = ColorLoss(coef=0.7)(a, tgt), color_loss(b, tgt, 0.7)
    torch.testing.assert_close(got, want, rtol=2e-5, atol=1e-7)
    got.backward(); want.backward()
    for k in a:
        torch.testing.assert_close(a[k].grad, b[k].grad, rtol=2e-5, atol=1e-9)
