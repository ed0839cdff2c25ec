"""Device-resident input pipeline: batches cut on the GPU from the byte-coded corpus must be exactly
what the reference's CharDataset.__getitem__ returns (tests/golden/ref_char_dataset.pt, minted from
mingpt/char_dataset.py:35-47), in the order its DataLoader + DistributedSampler would deliver them."""
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

import mingpt_distributed_b200 as m
from conftest import load_golden

pytestmark = pytest.mark.gpu


def _dataset():
    fx = load_golden("ref_char_dataset")
    ds = m.CharDataset(m.DataConfig(path=None, block_size=fx["block_size"], train_split=0.9, truncate=fx["truncate"]),
                       data=fx["corpus"])
    return fx, ds


def test_windows_equal_the_reference_items():
    fx, ds = _dataset()
    assert len(ds) == fx["length"] and ds.vocab_size == fx["vocab_size"] and ds.stoi == fx["stoi"]
    sampler = DistributedSampler(ds, num_replicas=1, rank=0, shuffle=False)
    loader = m.DeviceWindowLoader(ds, 100, "cuda", sampler)
    xs, ys = zip(*[(x.cpu(), y.cpu()) for x, y in loader])
    X, Y = torch.cat(xs), torch.cat(ys)
    assert X.shape == (len(ds), fx["block_size"]) and X.dtype == torch.int64
    for i, (rx, ry) in fx["items"].items():          # the reference's own __getitem__ outputs
        assert torch.equal(X[i], rx) and torch.equal(Y[i], ry), i
    for i in range(0, len(ds), 37):                  # and ours, which test_host pins to the same fixture
        a, b = ds[i]
        assert torch.equal(X[i], a) and torch.equal(Y[i], b)
    assert int(loader._err.item()) == 0


def test_shuffled_rank_shards_and_ragged_tail():
    _, ds = _dataset()
    seen = []
    for rank in range(2):
        sampler = DistributedSampler(ds, num_replicas=2, rank=rank)       # shuffle=True, seed=0: the reference's default
        want = list(iter(DistributedSampler(ds, num_replicas=2, rank=rank)))
        loader = m.DeviceWindowLoader(ds, 64, "cuda", sampler)
        got_x = torch.cat([x.cpu() for x, _ in loader])
        assert got_x.shape[0] == len(want)
        assert (len(want) % 64) != 0                                       # the last batch is ragged
        for j in (0, 1, 63, 64, len(want) - 1):
            assert torch.equal(got_x[j], ds[want[j]][0])
        seen += want
    assert set(seen) == set(range(len(ds)))


@pytest.mark.parametrize("T", [32, 33, 7])
def test_token_windows_kernel_edges(T):
    from mingpt_distributed_b200 import ops
    codes = torch.arange(0, 250, dtype=torch.uint8, device="cuda")
    starts = torch.tensor([0, 5, 250 - T - 1, 17], device="cuda")
    x = torch.full((4, T), -7, dtype=torch.int64, device="cuda"); y = torch.full_like(x, -7)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.token_windows(codes, starts, T, x, y, flag)
    for b, s in enumerate(starts.tolist()):
        assert x[b].tolist() == list(range(s, s + T)) and y[b].tolist() == list(range(s + 1, s + T + 1))
    assert flag.item() == 0
    bad = torch.tensor([250 - T, -1], device="cuda")               # one past the last valid start; negative
    xb = torch.full((2, T), -7, dtype=torch.int64, device="cuda"); yb = torch.full_like(xb, -7)
    ops.token_windows(codes, bad, T, xb, yb, flag)
    assert flag.item() == 3 and (xb == -7).all() and (yb == -7).all()


def test_gpttrainer_uses_the_device_loader_and_set_epoch_flag(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("LOCAL_RANK", "0"); monkeypatch.setenv("RANK", "0")
    _, ds = _dataset()
    gcfg = m.GPTConfig(model_type="gpt-nano", vocab_size=ds.vocab_size, block_size=ds.block_size, embed_drop=0, resid_drop=0, attn_drop=0)
    def trainer(**kw):
        model = m.GPT(gcfg, verbose=False)
        opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=3e-3))
        cfg = m.GPTTrainerConfig(max_epochs=2, batch_size=64, grad_norm_clip=1.0, dl_num_workers=0,
                                 snapshot_path=str(tmp_path / "none.pt"), save_every=10, **kw)
        return m.GPTTrainer(cfg, model, opt, ds)
    torch.manual_seed(0)
    a = trainer()
    assert isinstance(a.train_loader, m.DeviceWindowLoader)
    torch.manual_seed(0)
    b = trainer(device_loader=False)
    assert isinstance(b.train_loader, torch.utils.data.DataLoader)
    # same sampler order -> the two pipelines feed identical batches
    xa = [x.cpu() for x, _ in a.train_loader]
    xb = [x for x, _ in b.train_loader]
    assert len(xa) == len(xb) and all(torch.equal(p, q) for p, q in zip(xa, xb))
    a.train()
    assert a.graph_replays > 20
    c = trainer(shuffle_each_epoch=True)
    first = c.train_loader.window_starts()
    c.train_loader.sampler.set_epoch(1)
    assert not torch.equal(first, c.train_loader.window_starts())
