"""GPU: FusedAdam (one wgs_adam_step launch for all parameter tensors) against torch.optim.Adam run on the CPU
(plain PyTorch fp32 reference of the same op), with the reference's usage pattern: one param group per tensor,
eps 1e-15, the xyz learning rate rewritten every iteration, a parameter without gradient, and the in-place
state surgery densification performs (gaussian_model.py cat_tensors_to_optimizer / _prune_optimizer)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wildgs_slam_blackwell_b200.adam import FusedAdam  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = {"xyz": (5001, 3), "f_dc": (5001, 1, 3), "f_rest": (5001, 15, 3), "opacity": (5001, 1), "scaling": (5001, 3),
          "rotation": (5001, 4), "unused": (7,)}
LR = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3, "unused": 0.1}


def make(device, cls):
    g = torch.Generator().manual_seed(1)
    params = {k: torch.randn(*s, generator=g).to(device).requires_grad_(True) for k, s in SHAPES.items()}
    groups = [{"params": [params[k]], "lr": LR[k], "name": k} for k in SHAPES]
    return params, cls(groups, lr=0.0, eps=1e-15)


def test_matches_torch_adam_over_steps_with_state_surgery():
    pc, oc = make("cpu", torch.optim.Adam)
    pg, og = make(DEV, FusedAdam)
    gen = torch.Generator().manual_seed(2)
    for it in range(12):
        for k in SHAPES:
            if k == "unused" or (k == "f_rest" and it % 3 == 0):     # f_rest: no gradient on some steps
                pc[k].grad = pg[k].grad = None
                continue
            g = torch.randn(*pc[k].shape, generator=gen) * (10.0 ** (-it % 4))
            pc[k].grad, pg[k].grad = g.clone(), g.to(DEV)
        for o in (oc, og):                                            # update_learning_rate(iteration)
            o.param_groups[0]["lr"] = LR["xyz"] * (0.97 ** it)
        oc.step()
        og.step()
        if it == 5:                                                   # densify: append rows to xyz and its state
            for params, opt, dev in ((pc, oc, "cpu"), (pg, og, DEV)):
                grp = opt.param_groups[0]
                old = grp["params"][0]
                st = opt.state.pop(old)
                ext = torch.full((9, 3), 0.25, device=dev)
                st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)))
                st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)))
                new = torch.nn.Parameter(torch.cat((old.detach(), ext)).requires_grad_(True))
                grp["params"][0] = new
                opt.state[new] = st
                params["xyz"] = new
    for k in SHAPES:
        # each step moves a parameter by ~lr; fp32 rounding of 12 such steps (CPU and GPU contract differently)
        np.testing.assert_allclose(pg[k].detach().cpu().numpy(), pc[k].detach().numpy(), rtol=2e-5,
                                   atol=1e-3 * LR[k] * 12 + 1e-7, err_msg=k)
    np.testing.assert_array_equal(pg["unused"].detach().cpu().numpy(), pc["unused"].detach().numpy())
    for (kc, sc), (kg, sg) in zip(oc.state.items(), og.state.items()):
        m, v = sc["exp_avg"].numpy(), sc["exp_avg_sq"].numpy()      # moments: sums with cancellation -> atol on their scale
        np.testing.assert_allclose(sg["exp_avg"].cpu().numpy(), m, rtol=1e-5, atol=1e-6 * np.abs(m).max())
        np.testing.assert_allclose(sg["exp_avg_sq"].cpu().numpy(), v, rtol=1e-5, atol=1e-6 * np.abs(v).max())
        assert int(sg["step"]) == int(sc["step"])


def test_capturable_step_replays_from_a_cuda_graph():
    """capturable=True: the step count lives on the device and a tensor lr is read at run time, so ONE captured
    wgs_adam_step launch replayed 9 times equals 9 torch.optim.Adam steps (bias corrections and the lr schedule
    advance inside the graph)."""
    g = torch.Generator().manual_seed(4)
    w0 = torch.randn(3001, 3, generator=g)
    grads = [torch.randn(3001, 3, generator=g) for _ in range(9)]
    pc = w0.clone().requires_grad_(True)
    oc = torch.optim.Adam([{"params": [pc], "lr": 1e-2}], lr=0.0, eps=1e-15)
    pg = w0.to(DEV).requires_grad_(True)
    lr = torch.tensor(1e-2, device=DEV)
    og = FusedAdam([{"params": [pg], "lr": lr}], lr=0.0, eps=1e-15, capturable=True)
    gbuf = torch.zeros(3001, 3, device=DEV)
    pg.grad = gbuf
    graph = None
    for it, gr in enumerate(grads):
        oc.param_groups[0]["lr"] = 1e-2 * 0.9 ** it
        pc.grad = gr.clone()
        oc.step()
        gbuf.copy_(gr.to(DEV))
        lr.fill_(1e-2 * 0.9 ** it)
        if it < 2:
            og.step()                                    # eager steps create the state
        elif graph is None:
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                og.step()
            graph.replay()                               # capture does not execute
        else:
            graph.replay()
    st = og.state[pg]
    assert st["step"].is_cuda and float(st["step"]) == 9.0
    np.testing.assert_allclose(pg.detach().cpu().numpy(), pc.detach().numpy(), rtol=2e-5, atol=1e-3 * 1e-2 * 9)
    m, v = oc.state[pc]["exp_avg"].numpy(), oc.state[pc]["exp_avg_sq"].numpy()
    np.testing.assert_allclose(st["exp_avg"].cpu().numpy(), m, rtol=1e-5, atol=1e-6 * np.abs(m).max())
    np.testing.assert_allclose(st["exp_avg_sq"].cpu().numpy(), v, rtol=1e-5, atol=1e-6 * np.abs(v).max())


def test_errors():
    with pytest.raises(NotImplementedError):
        FusedAdam([torch.zeros(3, device=DEV, requires_grad=True)], amsgrad=True)
    p = torch.zeros(3, requires_grad=True)
    p.grad = torch.ones(3)
    with pytest.raises(RuntimeError, match="CUDA"):
        FusedAdam([p]).step()
