"""-m gpu: checkpoint compatibility with the reference's .tar format (utils.py:174-214,444-456):
a file written the way the reference writes it (nn.Module state_dicts + torch.optim.Adam state) resumes here, and a file
written here loads into torch.optim.Adam."""
import types

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests.gpu_util import make_model

pytestmark = pytest.mark.gpu


def _args(tmp_path):
    return types.SimpleNamespace(basedir=str(tmp_path), expname="exp", ft_path=None, no_reload=False, lrate=5e-4)


def _torch_twin(model):
    """fp32 nn.Linear stack with the reference's parameter names and shapes (nerf.py:79-94), for torch.optim.Adam."""
    twin = torch.nn.ParameterDict()
    ps = {k: torch.nn.Parameter(v.detach().clone()) for k, v in model.state_dict().items()}
    return ps


def test_reference_style_checkpoint_resumes_here(tmp_path):
    from nerf_shared_b200 import utils
    args = _args(tmp_path)
    mc, mf = make_model(11), make_model(12)
    # the reference side: torch.optim.Adam over coarse then fine parameters (utils.py:163-172), two steps taken
    pc, pf = _torch_twin(mc), _torch_twin(mf)
    ref_params = list(pc.values()) + list(pf.values())
    ref_opt = torch.optim.Adam(ref_params, lr=5e-4, betas=(0.9, 0.999))
    g = torch.Generator().manual_seed(0)
    grads = [[torch.randn(p.shape, generator=g).cuda() * 1e-3 for p in ref_params] for _ in range(3)]
    for step in range(2):
        for p, gr in zip(ref_params, grads[step]):
            p.grad = gr.clone()
        ref_opt.step()
    import os
    os.makedirs(os.path.join(args.basedir, args.expname), exist_ok=True)
    torch.save({"global_step": 2, "coarse_model_state_dict": {k: v.detach() for k, v in pc.items()},
                "fine_model_state_dict": {k: v.detach() for k, v in pf.items()},
                "optimizer_state_dict": ref_opt.state_dict()}, os.path.join(args.basedir, args.expname, "000002.tar"))
    # this side resumes and takes the third step; torch takes it too
    opt = utils.get_optimizer(mc, mf, args)
    start = utils.load_checkpoint(mc, mf, opt, args, b_load_ckpnt_as_trainable=True)
    assert start == 2
    ours = list(mc.parameters()) + list(mf.parameters())
    for p, q in zip(ours, ref_params):
        assert torch.equal(p.detach(), q.detach())
    opt.zero_grad()
    for p, gr in zip(ours, grads[2]):
        p.grad.copy_(gr)
    opt.step()
    for p, gr in zip(ref_params, grads[2]):
        p.grad = gr.clone()
    ref_opt.step()
    for p, q in zip(ours, ref_params):
        np.testing.assert_allclose(p.detach().cpu().numpy(), q.detach().cpu().numpy(), rtol=0, atol=2e-7)


def test_checkpoint_written_here_loads_into_torch_adam(tmp_path):
    from nerf_shared_b200 import utils
    args = _args(tmp_path)
    mc, mf = make_model(11), make_model(12)
    opt = utils.get_optimizer(mc, mf, args)
    ours = list(mc.parameters()) + list(mf.parameters())
    g = torch.Generator().manual_seed(1)
    for _ in range(2):
        opt.zero_grad()
        for p in ours:
            p.grad.copy_(torch.randn(p.shape, generator=g).cuda() * 1e-3)
        opt.step()
    utils.save_checkpoints(args, mc, mf, opt, global_step=2, i=2)
    import os
    ckpt = torch.load(os.path.join(args.basedir, args.expname, "000002.tar"), map_location="cuda")
    assert set(ckpt) == {"global_step", "coarse_model_state_dict", "fine_model_state_dict", "optimizer_state_dict"}
    assert list(ckpt["coarse_model_state_dict"]) == list(O.make_params(11))         # the reference's keys, in order
    pc = {k: torch.nn.Parameter(v.clone()) for k, v in ckpt["coarse_model_state_dict"].items()}
    pf = {k: torch.nn.Parameter(v.clone()) for k, v in ckpt["fine_model_state_dict"].items()}
    ref_params = list(pc.values()) + list(pf.values())
    ref_opt = torch.optim.Adam(ref_params, lr=5e-4, betas=(0.9, 0.999))
    ref_opt.load_state_dict(ckpt["optimizer_state_dict"])
    # one more identical step on both sides
    grads = [torch.randn(p.shape, generator=g).cuda() * 1e-3 for p in ours]
    opt.zero_grad()
    for p, gr in zip(ours, grads):
        p.grad.copy_(gr)
    opt.step()
    for p, gr in zip(ref_params, grads):
        p.grad = gr.clone()
    ref_opt.step()
    for p, q in zip(ours, ref_params):
        np.testing.assert_allclose(p.detach().cpu().numpy(), q.detach().cpu().numpy(), rtol=0, atol=2e-7)
