"""The fused LARC + SGD-momentum step (ntx_larc_sgd_step through simclr_b200.larc.LARC) against the fp64 oracle."""
import numpy as np
import pytest
import torch

from oracle import larc_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from simclr_b200 import _cabi
    _cabi.load()
    return torch.device("cuda:0")


@pytest.mark.parametrize("clip", [False, True])
def test_fused_step_matches_oracle_over_several_steps(clip, dev):
    """Ragged tensor sizes (unaligned tails, one element, several chunks), a weight-decay group and a no-decay group
    (exclude_from_wt_decay, main.py:18-36), an all-zero parameter and an all-zero gradient (left un-adapted like
    apex does), momentum carried over four steps with a changing learning rate.  fp32 kernel vs fp64 oracle:
    1e-6 of the largest parameter."""
    from simclr_b200.larc import LARC
    rng = np.random.default_rng(5)
    shapes = [(64, 3, 3, 3), (64,), (70000,), (1,), (513, 129), (8191,), (3, 8192)]
    decay_group, plain_group = [0, 2, 4, 6], [1, 3, 5]
    mom, wd, trust, eps = 0.9, 1e-4, 0.001, 1e-8
    init = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    init[5][:] = 0.0
    tp = [torch.nn.Parameter(torch.tensor(x, device=dev)) for x in init]
    sgd = torch.optim.SGD([{"params": [tp[i] for i in decay_group], "weight_decay": wd},
                           {"params": [tp[i] for i in plain_group], "weight_decay": 0.0}],
                          lr=0.5, momentum=mom, nesterov=False, weight_decay=0.0)
    opt = LARC(optimizer=sgd, trust_coefficient=trust, clip=clip, eps=eps)
    params = [x.astype(np.float64) for x in init]
    bufs = [None] * len(shapes)
    for step in range(4):
        lr = 0.5 / (step + 1)
        for group in opt.param_groups:
            group["lr"] = lr
        grads = [rng.standard_normal(s).astype(np.float32) for s in shapes]
        if step == 2:
            grads[1][:] = 0.0
        opt.zero_grad()
        for p, g in zip(tp, grads):
            p.grad = torch.tensor(g, device=dev)
        opt.step()
        for idx, w in ((decay_group, wd), (plain_group, 0.0)):
            new_p, new_b = larc_oracle.larc_sgd_step([params[i] for i in idx], [grads[i] for i in idx],
                                                     [bufs[i] for i in idx], lr, mom, w, trust, clip, eps)
            for i, a, b in zip(idx, new_p, new_b):
                params[i], bufs[i] = a, b
        for i, (p, want) in enumerate(zip(tp, params)):
            got = p.detach().cpu().numpy().astype(np.float64)
            assert np.abs(got - want).max() <= 1e-6 * max(1.0, np.abs(want).max()), (step, i)
            mb = sgd.state[p]["momentum_buffer"].cpu().numpy().astype(np.float64)
            assert np.abs(mb - bufs[i]).max() <= 1e-6 * max(1.0, np.abs(bufs[i]).max()), (step, i)
    # state round-trips with plain torch.optim.SGD
    plain = torch.optim.SGD([{"params": [tp[i] for i in decay_group]}, {"params": [tp[i] for i in plain_group]}],
                            lr=0.1, momentum=mom)
    plain.load_state_dict(opt.state_dict())
    assert torch.equal(plain.state[tp[0]]["momentum_buffer"], sgd.state[tp[0]]["momentum_buffer"])


def test_fused_step_is_deterministic_and_matches_torch_on_a_resnet(dev):
    """ResNet-18-sized parameter set (the reference's base_cnn): two identical runs are bit-identical, and the
    result agrees with torch.optim.SGD fed LARC-edited gradients (torch ops in fp32) to fp32 rounding."""
    import torchvision
    from simclr_b200.larc import LARC
    torch.manual_seed(0)
    nets = [torchvision.models.resnet18(num_classes=128).to(dev) for _ in range(3)]
    for n in nets[1:]:
        n.load_state_dict(nets[0].state_dict())
    grads = [torch.randn_like(p) for p in nets[0].parameters()]

    def groups(net):
        named = list(net.named_parameters())
        skip = ("bias", "bn")
        return [{"params": [p for n, p in named if not any(s in n for s in skip)], "weight_decay": 1e-4},
                {"params": [p for n, p in named if any(s in n for s in skip)], "weight_decay": 0.0}]

    outs = []
    for net in nets[:2]:
        opt = LARC(torch.optim.SGD(groups(net), lr=0.3, momentum=0.9), trust_coefficient=0.001, clip=False)
        for _ in range(2):
            for p, g in zip(net.parameters(), grads):
                p.grad = g.clone()
            opt.step()
        outs.append([p.detach().clone() for p in net.parameters()])
    assert all(torch.equal(a, b) for a, b in zip(*outs))
    ref = torch.optim.SGD(groups(nets[2]), lr=0.3, momentum=0.9)
    for _ in range(2):
        for p, g in zip(nets[2].parameters(), grads):
            p.grad = g.clone()
        with torch.no_grad():
            for group in ref.param_groups:
                wd = group["weight_decay"]
                for p in group["params"]:
                    pn, gn = torch.norm(p), torch.norm(p.grad)
                    if pn != 0 and gn != 0:
                        p.grad.add_(p, alpha=wd).mul_(0.001 * pn / (gn + pn * wd + 1e-8))
                group["weight_decay"] = 0.0
            ref.step()
            for group, wd in zip(ref.param_groups, (1e-4, 0.0)):
                group["weight_decay"] = wd
    for a, b in zip(outs[0], nets[2].parameters()):
        assert float((a - b.detach()).abs().max()) <= 2e-6 * max(1.0, float(b.detach().abs().max()))
