"""ORACLE -- test infrastructure only. Generates tests/golden/ref_*.pt by running the repaired REFERENCE
(oracle/reference_loader.py) on seeded inputs in the build container (the only place /root/reference
exists). Run:  python -m oracle.make_golden     (deterministic: CPU, fp32, fixed seeds)

Each fixture holds: config, torch version, the reference's initial state_dict at torch.manual_seed(0),
a token batch with ~10 % ignored targets, the reference's logits / loss / every parameter gradient on it,
a 10-step loss trajectory of the restated GPTTrainer._run_batch (mingpt/trainer.py:118-133: forward,
zero_grad, backward, clip_grad_norm 1.0, AdamW from the reference's own create_optimizer), the state
after those steps, and the reference's greedy generate() continuation.
"""
from __future__ import annotations

import os

import torch

from . import reference_loader as rl

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = {
    # name: (n_layer, n_head, n_embed, vocab, block, B, T)
    "ref_nano": (3, 3, 48, 65, 128, 4, 64),      # gpt-nano preset sizes (mingpt/model.py:292-293), d = 16
    "ref_d32": (2, 2, 64, 97, 64, 3, 40),        # d = 32 (gpt-mini's head size), ragged T and vocab
    "ref_d64": (1, 2, 128, 130, 96, 2, 96),      # d = 64 (gpt2 family head size)
}


def make(name, n_layer, n_head, n_embed, vocab, block, B, T):
    mod, model = rl.build_reference_gpt(n_layer, n_head, n_embed, vocab, block, seed=0)
    sd0 = {k: v.clone() for k, v in model.state_dict().items() if not k.endswith(".attn.mask")}
    g = torch.Generator().manual_seed(1234)
    tok = torch.randint(0, vocab, (B, T + 1), generator=g)
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].clone()
    drop = torch.rand(B, T, generator=g) < 0.1
    y[drop] = -1
    logits, loss = model(x, y)
    model.zero_grad(set_to_none=True)
    loss.backward()
    grads = {n: p.grad.clone() for n, p in model.named_parameters()}
    # 10 optimizer steps, mingpt/trainer.py:118-133
    opt = mod.create_optimizer(model, mod.OptimizerConfig())
    traj = []
    for step in range(10):
        gs = torch.Generator().manual_seed(100 + step)
        tk = torch.randint(0, vocab, (B, T + 1), generator=gs)
        _, l = model(tk[:, :-1].contiguous(), tk[:, 1:].contiguous())
        opt.zero_grad(set_to_none=True)
        l.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        traj.append(l.item())
    sd10 = {k: v.clone() for k, v in model.state_dict().items() if not k.endswith(".attn.mask")}
    prompt = tok[:, :8].contiguous()
    gen = model.generate(prompt, max_new_tokens=12)
    fixture = dict(
        config=dict(n_layer=n_layer, n_head=n_head, n_embed=n_embed, vocab_size=vocab, block_size=block),
        torch_version=torch.__version__, state_dict=sd0, x=x, y=y, logits=logits.detach(), loss=loss.item(),
        grads=grads, traj=traj, state_dict_after=sd10, prompt=prompt, generated=gen,
    )
    os.makedirs(OUT, exist_ok=True)
    torch.save(fixture, os.path.join(OUT, name + ".pt"))
    print(name, "loss", loss.item(), "traj", [round(t, 4) for t in traj])


def make_char_dataset_fixture():
    """CharDataset of the reference (mingpt/char_dataset.py) on a seeded byte corpus."""
    import importlib.util, tempfile, contextlib, io
    spec = importlib.util.spec_from_file_location("_reference_char_dataset", os.path.join(rl.REFERENCE_ROOT, "mingpt", "char_dataset.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    g = torch.Generator().manual_seed(7)
    alphabet = bytes(range(97, 123)) + b" .,\n"
    corpus = bytes(alphabet[i] for i in torch.randint(0, len(alphabet), (4096,), generator=g).tolist())
    with tempfile.NamedTemporaryFile(suffix=".txt", delete=False) as f:
        f.write(corpus)
        path = f.name
    with contextlib.redirect_stdout(io.StringIO()):
        ds = mod.CharDataset(mod.DataConfig(path=path, block_size=32, train_split=0.9, truncate=0.5))
    items = {i: ds[i] for i in (0, 1, 17, len(ds) - 1)}
    torch.save(dict(corpus=corpus, block_size=32, truncate=0.5, length=len(ds), vocab_size=ds.vocab_size,
                    stoi=ds.stoi, items=items), os.path.join(OUT, "ref_char_dataset.pt"))
    os.unlink(path)
    print("char dataset", len(ds), ds.vocab_size)


if __name__ == "__main__":
    assert rl.available(), "needs /root/reference"
    torch.set_num_threads(1)
    for name, args in CASES.items():
        make(name, *args)
    make_char_dataset_fixture()
