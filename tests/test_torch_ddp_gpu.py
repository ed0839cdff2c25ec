"""The reference wraps its model in torch.nn.parallel.DistributedDataParallel (mingpt/trainer.py:71) and then runs
forward / zero_grad / backward / clip_grad_norm / optimizer.step (:118-133). That exact sequence must keep working
on the fused GPT: DDP's reducer hooks hang on autograd's AccumulateGrad nodes, so the model switches to returning
its parameter gradients through autograd when it finds itself called from DDP. Single process, NCCL world size 1
(the hooks, bucket copies and the all-reduce all run; with one rank the average is the identity)."""
import os

import pytest
import torch
import torch.distributed as dist

import mingpt_distributed_b200 as m
from conftest import report

pytestmark = pytest.mark.gpu


@pytest.fixture
def nccl_world1():
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29631")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield
    dist.destroy_process_group()


def _make(seed=0):
    torch.manual_seed(seed)
    cfg = m.GPTConfig(model_type=None, n_layer=2, n_head=2, n_embed=128, vocab_size=211, block_size=64, embed_drop=0,
                      resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda()
    opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=1e-3))
    return model, opt


def _batch(step):
    g = torch.Generator().manual_seed(70 + step)
    tok = torch.randint(0, 211, (4, 65), generator=g)
    return tok[:, :-1].contiguous().cuda(), tok[:, 1:].contiguous().cuda()


def test_reference_run_batch_sequence_under_torch_ddp(nccl_world1):
    from torch.nn.parallel import DistributedDataParallel as DDP
    plain, opt_p = _make()
    wrapped, opt_w = _make()
    fired = []
    with pytest.warns(UserWarning, match="DistributedDataParallel"):
        ddp = DDP(wrapped, device_ids=[0])

        def counting_hook(state, bucket):
            fired.append(bucket.buffer().numel())
            return dist.algorithms.ddp_comm_hooks.default_hooks.allreduce_hook(state, bucket)

        import torch.distributed.algorithms.ddp_comm_hooks.default_hooks  # noqa: F401
        ddp.register_comm_hook(None, counting_hook)
        losses_w, losses_p = [], []
        for step in range(4):
            x, y = _batch(step)
            # mingpt/trainer.py:122-132, as written
            _, loss = ddp(x, y)
            opt_w.zero_grad(set_to_none=True)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(ddp.parameters(), 1.0)
            opt_w.step()
            losses_w.append(loss.item())
            _, lp = plain(x, y)
            opt_p.zero_grad(set_to_none=True)
            lp.backward()
            torch.nn.utils.clip_grad_norm_(plain.parameters(), 1.0)
            opt_p.step()
            losses_p.append(lp.item())
    assert wrapped.autograd_param_grads and not plain.autograd_param_grads
    n_params = sum(p.numel() for p in wrapped.parameters())
    assert sum(fired) == 4 * n_params, "DDP's reducer must have seen every gradient of every step"
    dev = max(abs(a - b) / abs(b) for a, b in zip(losses_w, losses_p))
    worst = max(((a - b).norm() / (b.norm() + 1e-12)).item()
                for (_, a), (_, b) in zip(wrapped.named_parameters(), plain.named_parameters()) if a.dim() == 2)
    report("torch_ddp_vs_plain", max_loss_dev=dev, worst_param_rel=worst, buckets=len(fired))
    assert dev < 2e-3 and worst < 5e-3


def test_autograd_visible_gradients_accumulate_like_plain():
    """GPT.autograd_param_grads without DDP: two backward passes accumulate (AccumulateGrad adds the second)"""
    a, _ = _make(1)
    b, _ = _make(1)
    b.autograd_param_grads = True
    x, y = _batch(0)
    x2, y2 = _batch(1)
    for mod in (a, b):
        _, l1 = mod(x, y); l1.backward()
        _, l2 = mod(x2, y2); (l2 * 0.5).backward()
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(pa.grad, pb.grad, rtol=2e-2, atol=1e-5 * (1 + pa.grad.abs().max().item())), n
