"""FusedAdam (csrc/adam.cu) against torch.optim.Adam -- the optimizer the reference's drivers
build (train.py:331-335, cross_validation.py:93-97) -- on the same parameters and gradients."""
import pytest
import torch


def _params(seed, device):
    g = torch.Generator().manual_seed(seed)
    shapes = [(64, 1, 7, 7), (64,), (128, 64, 3, 3), (5,), (1, 16, 1, 1), (3, 100003), (1,), (256, 128, 3, 3)]
    return [torch.nn.Parameter(torch.randn(s, generator=g).to(device)) for s in shapes]


@pytest.mark.gpu
@pytest.mark.parametrize("wd", [0.0, 1e-5, 1e-2])
def test_fused_adam_matches_torch_adam(cuda, wd):
    from b2seg.optim import FusedAdam
    ours, ref = _params(3, cuda), _params(3, cuda)
    opt_o = FusedAdam(ours, lr=1e-3, weight_decay=wd)
    opt_r = torch.optim.Adam(ref, lr=1e-3, weight_decay=wd, foreach=False)
    sched_o = torch.optim.lr_scheduler.ReduceLROnPlateau(opt_o, mode='max', factor=0.5, patience=0)
    sched_r = torch.optim.lr_scheduler.ReduceLROnPlateau(opt_r, mode='max', factor=0.5, patience=0)
    g = torch.Generator().manual_seed(9)
    for it in range(12):
        for po, pr in zip(ours, ref):
            if it == 4 and po.numel() == 5:
                po.grad = pr.grad = None  # a parameter that skips a step keeps its own step count
                continue
            gr = (torch.randn(po.shape, generator=g) * (0.1 + it)).to(cuda)
            po.grad, pr.grad = gr.clone(), gr.clone()
        opt_o.step()
        opt_r.step()
        if it % 3 == 2:  # the plateau scheduler halves lr through param_groups
            sched_o.step(0.0)
            sched_r.step(0.0)
    assert opt_o.param_groups[0]["lr"] == opt_r.param_groups[0]["lr"] < 1e-3
    for po, pr in zip(ours, ref):
        torch.testing.assert_close(po, pr, rtol=2e-5, atol=2e-7)
        so, sr = opt_o.state[po], opt_r.state[pr]
        assert float(so["step"]) == float(sr["step"])
        # moments are sums of gradients of magnitude ~10: one fp32 ulp of a term is ~1e-6
        torch.testing.assert_close(so["exp_avg"], sr["exp_avg"], rtol=1e-4, atol=3e-6)
        torch.testing.assert_close(so["exp_avg_sq"], sr["exp_avg_sq"], rtol=1e-4, atol=1e-5)
    # state_dict interchange: the stock optimizer continues from our state and vice versa
    opt_r2 = torch.optim.Adam(ref, lr=1e-3, weight_decay=wd, foreach=False)
    opt_r2.load_state_dict(opt_o.state_dict())
    opt_o2 = FusedAdam(ours, lr=1e-3, weight_decay=wd)
    opt_o2.load_state_dict(opt_r.state_dict())
    for po, pr in zip(ours, ref):
        gr = torch.ones_like(po)
        po.grad, pr.grad = gr.clone(), gr.clone()
    opt_o2.step()
    opt_r2.step()
    for po, pr in zip(ours, ref):
        torch.testing.assert_close(po, pr, rtol=2e-5, atol=2e-7)


def test_fused_adam_is_a_torch_adam_and_has_no_cpu_path():
    from b2seg.optim import FusedAdam
    from b2seg._lib import B2SegError
    p = torch.nn.Parameter(torch.zeros(4))
    opt = FusedAdam([p], lr=1e-3, weight_decay=1e-5)
    assert isinstance(opt, torch.optim.Adam)
    assert opt.param_groups[0]["weight_decay"] == 1e-5 and opt.param_groups[0]["betas"] == (0.9, 0.999)
    opt.step()  # no gradients yet: nothing to do
    p.grad = torch.ones(4)
    with pytest.raises(B2SegError):
        opt.step()
    with pytest.raises(B2SegError):
        FusedAdam([torch.nn.Parameter(torch.zeros(2))], amsgrad=True).step()
