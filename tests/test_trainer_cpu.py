"""Public Trainer API on the host backend (no GPU): prefetch + step, periodic checkpoints."""
import os

import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss


def _ctx():
    return psd.Context(0, 1, 0, torch.device("cpu"), None)


def test_prefetch_then_step_equals_step_with_batch():
    cfg = MLPConfig(d_in=16, hidden=(32,), d_out=4)
    gen = torch.Generator().manual_seed(0)
    batches = [(torch.randn(8, 16, generator=gen), torch.randn(8, 4, generator=gen)) for _ in range(4)]
    runs = []
    for pre in (False, True):
        torch.manual_seed(1)
        tr = psd.Trainer(MLP(cfg), mse_loss, _ctx(), psd.OptimConfig(kind="sgd", lr=0.1), bucket_mb=0.001)
        if pre:
            losses = []
            tr.prefetch(*batches[0])
            for i in range(len(batches)):
                h = tr.step()
                if i + 1 < len(batches):
                    tr.prefetch(*batches[i + 1])
                losses.append(h.item())
        else:
            losses = [tr.step(*b).item() for b in batches]
        runs.append(losses)
    assert runs[0] == runs[1]


def test_periodic_checkpoints_are_written_atomically(tmp_path):
    cfg = MLPConfig(d_in=16, hidden=(32,), d_out=4)
    path = str(tmp_path / "run.ckpt")
    torch.manual_seed(2)
    tr = psd.Trainer(MLP(cfg), mse_loss, _ctx(), psd.OptimConfig(kind="adamw", lr=0.01),
                     bucket_mb=0.001, checkpoint_every=3, checkpoint_path=path)
    x, y = torch.randn(8, 16), torch.randn(8, 4)
    for i in range(7):
        tr.step(x, y)
        assert os.path.exists(path) == (i >= 2)
        assert not os.path.exists(path + ".tmp")
    epoch, it = tr.load_checkpoint(path)
    assert (epoch, it) == (2, 5)  # written after step 6 = iteration index 5
    assert os.path.exists(path + ".psb")  # AdamW moments ride in the sidecar


def test_learning_rate_schedule_through_set_lr():
    """ps.set_lr between steps == the same schedule applied by torch.optim.SGD."""
    cfg = MLPConfig(d_in=16, hidden=(32,), d_out=4)
    torch.manual_seed(3)
    model = MLP(cfg)
    ref = MLP(cfg)
    ref.load_state_dict(model.state_dict())
    tr = psd.Trainer(model, mse_loss, _ctx(), psd.OptimConfig(kind="sgd", lr=0.2), bucket_mb=0.001)
    opt = torch.optim.SGD(ref.parameters(), lr=0.2)
    gen = torch.Generator().manual_seed(4)
    for step in range(5):
        lr = 0.2 * (0.5 ** step)
        tr.ps.set_lr(lr)
        for g in opt.param_groups:
            g["lr"] = lr
        x, y = torch.randn(8, 16, generator=gen), torch.randn(8, 4, generator=gen)
        tr.step(x, y)
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        opt.step()
    for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
        assert torch.allclose(a, b, atol=1e-6), n
