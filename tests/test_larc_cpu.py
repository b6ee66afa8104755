"""The LARC + SGD oracle against the real torch.optim.SGD (CPU), and the wrapper's argument checks."""
import numpy as np
import pytest
import torch

from oracle import larc_oracle


def _apex_style_step(opt, trust, clip, eps):
    """LARC.step as apex publishes it, with torch ops, around the REAL torch.optim.SGD (apex is not installable)."""
    with torch.no_grad():
        decays = []
        for group in opt.param_groups:
            wd = group.get("weight_decay", 0)
            decays.append(wd)
            group["weight_decay"] = 0
            for p in group["params"]:
                if p.grad is None:
                    continue
                pn, gn = torch.norm(p.data), torch.norm(p.grad.data)
                if pn != 0 and gn != 0:
                    a = trust * pn / (gn + pn * wd + eps)
                    if clip:
                        a = min(a / group["lr"], 1)
                    p.grad.data += wd * p.data
                    p.grad.data *= a
    opt.step()
    for group, wd in zip(opt.param_groups, decays):
        group["weight_decay"] = wd


@pytest.mark.parametrize("clip", [False, True])
def test_oracle_matches_torch_sgd_with_larc_edits(clip):
    rng = np.random.default_rng(0)
    shapes = [(7, 5), (33,), (4, 3, 3, 3), (1,)]
    lr, mom, wd, trust, eps = 0.3, 0.9, 1e-3, 0.001, 1e-8
    init = [rng.standard_normal(s) for s in shapes]
    init[1][:] = 0.0                                       # a zero parameter tensor: left un-adapted, no decay
    tp = [torch.nn.Parameter(torch.tensor(x, dtype=torch.float64)) for x in init]
    opt = torch.optim.SGD([{"params": tp[:2], "weight_decay": wd}, {"params": tp[2:], "weight_decay": 0.0}],
                          lr=lr, momentum=mom, nesterov=False, weight_decay=0.0)
    params = [x.copy() for x in init]
    bufs = [None] * len(shapes)
    for step in range(3):
        grads = [rng.standard_normal(s) for s in shapes]
        if step == 1:
            grads[0][:] = 0.0                              # a zero gradient: same
        for p, g in zip(tp, grads):
            p.grad = torch.tensor(g, dtype=torch.float64)
        _apex_style_step(opt, trust, clip, eps)
        params[:2], bufs[:2] = larc_oracle.larc_sgd_step(params[:2], grads[:2], bufs[:2], lr, mom, wd, trust, clip, eps)
        params[2:], bufs[2:] = larc_oracle.larc_sgd_step(params[2:], grads[2:], bufs[2:], lr, mom, 0.0, trust, clip, eps)
        for p, want in zip(tp, params):
            assert np.abs(p.detach().numpy() - want).max() <= 1e-12 * max(1.0, np.abs(want).max())


def test_wrapper_rejects_what_it_does_not_fuse():
    from simclr_b200.larc import LARC
    p = torch.nn.Parameter(torch.zeros(3))
    with pytest.raises(TypeError):
        LARC(torch.optim.Adam([p]))
    opt = LARC(torch.optim.SGD([p], lr=0.1, momentum=0.9), trust_coefficient=0.001, clip=False)
    assert opt.optim.param_groups is opt.param_groups and opt.state_dict()["param_groups"][0]["lr"] == 0.1
    p.grad = torch.ones(3)
    with pytest.raises(TypeError, match="no CPU path"):
        opt.step()
    nest = LARC(torch.optim.SGD([p], lr=0.1, momentum=0.9, nesterov=True))
    with pytest.raises(ValueError, match="nesterov"):
        nest.step()
