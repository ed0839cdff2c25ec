"""Host-side logic and the drop-in API, no GPU: construction / initialisation parity, state_dict layout,
optimizer groups, config loading, dataset semantics, bucket planning, and that the C-ABI library loads
and exports every symbol include/mingpt_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest
import torch

import mingpt_distributed_b200 as m
from conftest import ROOT, load_golden
from mingpt_distributed_b200 import _lib

NANO = dict(model_type=None, n_layer=3, n_head=3, n_embed=48, vocab_size=65, block_size=128, embed_drop=0.0,
            resid_drop=0.0, attn_drop=0.0)


def test_reference_variant_init_is_bit_identical_to_reference():
    fx = load_golden("ref_nano")
    torch.manual_seed(0)
    g = m.GPT(m.GPTConfig(variant="reference", **NANO), verbose=False)
    sd = g.state_dict()
    assert list(sd.keys()) == list(fx["state_dict"].keys())
    for k, v in fx["state_dict"].items():
        assert torch.equal(sd[k], v), k


def test_parameter_counts():
    # as-written / mingpt-variant parameter counts
    ref = m.GPT(m.GPTConfig(variant="reference", **NANO), verbose=False)
    mg = m.GPT(m.GPTConfig(variant="mingpt", **NANO), verbose=False)
    assert sum(p.numel() for p in ref.parameters()) == 104352
    assert sum(p.numel() for p in mg.parameters()) == 97296


def test_presets_and_aliases():
    g = m.GPT(m.GPTConfig(model_type="gpt-mini", vocab_size=65, block_size=256), verbose=False)
    assert (g.n_layer, g.n_head, g.n_embed) == (6, 6, 192)
    g = m.GPT(m.GPTConfig(model_type=None, n_layer=2, n_head=2, n_embd=32, vocab_size=11, block_size=8), verbose=False)
    assert g.n_embed == 32
    with pytest.raises(ValueError):
        m.GPT(m.GPTConfig(model_type="nope", vocab_size=11, block_size=8), verbose=False)


def test_readme_api_spelling():
    cfg = m.GPT.get_default_config()
    cfg.model_type, cfg.vocab_size, cfg.block_size = "gpt-nano", 65, 128
    g = m.GPT(cfg, verbose=False)
    assert g.variant == "mingpt" and "lm_head.weight" in g.state_dict()
    tc = m.Trainer.get_default_config()
    opt = g.configure_optimizers(tc)
    assert isinstance(opt, m.FusedAdamW) and opt.defaults["betas"] == (0.9, 0.95)


def test_flat_storage_views_and_state_dict_roundtrip():
    g = m.GPT(m.GPTConfig(variant="reference", **NANO), verbose=False)
    base = g._flat.data_ptr()
    for n, p in g.named_parameters():
        o, k = g._spans[n]
        assert p.data_ptr() == base + 4 * o and p.numel() == k and o % 64 == 0
    sd = {k: v.clone() for k, v in g.state_dict().items()}
    # a reference snapshot also carries per-layer mask buffers (mingpt/model.py:142-145): must load
    sd["blocks.0.attn.mask"] = torch.ones(1, 1, 128, 128)
    g2 = m.GPT(m.GPTConfig(variant="reference", **NANO), verbose=False)
    g2.load_state_dict(sd)
    for k, v in g2.state_dict().items():
        assert torch.equal(v, sd[k])
    spans = g._layer_spans
    assert len(spans) == g.n_layer + 2 and spans[0][0] == 0 and spans[-1][1] == g._total
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def test_param_groups_match_reference_rules():
    for variant, want in (("reference", (16, 31)), ("mingpt", (13, 28))):
        g = m.GPT(m.GPTConfig(variant=variant, **NANO), verbose=False)
        d, nd = m.param_groups(g)
        assert (len(d), len(nd)) == want
        opt = m.create_optimizer(g, m.OptimizerConfig())
        assert [gr["weight_decay"] for gr in opt.param_groups] == [0.1, 0.0]
        assert opt.defaults["lr"] == 3e-4 and opt.defaults["eps"] == 1e-8
        names = {id(p): n for n, p in g.named_parameters()}
        assert [names[id(p)] for p in opt.param_groups[0]["params"]] == d   # sorted by name, model.py:107-116


def test_optimizer_state_dict_layout_matches_torch_adamw():
    g = m.GPT(m.GPTConfig(variant="mingpt", **NANO), verbose=False)
    ours = m.create_optimizer(g, m.OptimizerConfig()).state_dict()
    d, nd = m.param_groups(g)
    pd = dict(g.named_parameters())
    theirs = torch.optim.AdamW([{"params": [pd[n] for n in d], "weight_decay": 0.1},
                                {"params": [pd[n] for n in nd], "weight_decay": 0.0}], lr=3e-4, betas=(0.9, 0.95)).state_dict()
    for a, b in zip(ours["param_groups"], theirs["param_groups"]):
        assert a["params"] == b["params"]
        for k in ("lr", "betas", "eps", "weight_decay"):
            assert a[k] == b[k]


def test_cpu_forward_fails_loudly():
    g = m.GPT(m.GPTConfig(variant="mingpt", **NANO), verbose=False)
    with pytest.raises(RuntimeError, match="CUDA"):
        g(torch.zeros(1, 8, dtype=torch.long))


def test_config_loader_accepts_reference_yaml(tmp_path):
    ref_yaml = os.path.join(ROOT, "mingpt_distributed_b200", "gpt2_config.yaml")
    gc, oc, dc, tc = m.load_config(ref_yaml, ["gpt_config.n_layer=2", "trainer_config.batch_size=8"])
    assert gc.n_layer == 2 and gc.n_embed == 512 and gc.n_head == 8     # n_embd in the YAML
    assert oc.learning_rate == 3e-4 and tc.batch_size == 8 and dc.block_size == 128
    # the reference's own file (hydra section ignored, s3 paths kept as strings)
    p = tmp_path / "ref.yaml"
    p.write_text("gpt_config:\n  n_layer: 8\n  n_head: 8\n  n_embd: 512\noptimizer_config:\n  weight_decay: 0.1\n"
                 "  learning_rate: 0.0003\ndata_config:\n  path: s3://bucket\n  block_size: 128\n  truncate: 0.05\n"
                 "trainer_config:\n  max_epochs: 10\n  batch_size: 64\n  dl_num_workers: 4\n  grad_norm_clip: 1.0\n"
                 "  snapshot_path: s3://bucket/gpt_snapshot.pt\n  save_every: 3\nhydra:\n  run:\n    dir: ./\n")
    gc, oc, dc, tc = m.load_config(str(p))
    assert gc.n_embed == 512 and tc.save_every == 3 and dc.path == "s3://bucket"
    with pytest.raises(TypeError):
        m.load_config(str(p), ["gpt_config.bogus=1"])


def test_char_dataset_matches_reference_fixture(tmp_path):
    fx = load_golden("ref_char_dataset")
    path = tmp_path / "corpus.txt"
    path.write_bytes(fx["corpus"])
    ds = m.CharDataset(m.DataConfig(path=str(path), block_size=fx["block_size"], train_split=0.9, truncate=fx["truncate"]))
    assert len(ds) == fx["length"] and ds.vocab_size == fx["vocab_size"] and ds.stoi == fx["stoi"]
    for i, (x, y) in fx["items"].items():
        a, b = ds[i]
        assert torch.equal(a, x) and torch.equal(b, y) and a.dtype == torch.int64
    empty = m.CharDataset(m.DataConfig(path=str(path), block_size=4, truncate=0.0))
    assert empty.vocab_size == 0 and empty.__len__() == -4     # the reference's arithmetic on an empty corpus


def test_plan_buckets_reverse_order_and_coverage():
    spans = [(0, 100), (100, 300), (300, 500), (500, 700), (700, 1000)]
    b = m.plan_buckets(spans, cap_elems=350)
    assert b[0][1] == 1000 and b[-1][0] == 0
    assert [x[2] for x in b] == sorted([x[2] for x in b], reverse=True)
    covered = sorted((s, e) for s, e, _ in b)
    assert covered[0][0] == 0 and all(a[1] == c[0] for a, c in zip(covered, covered[1:])) and covered[-1][1] == 1000
    assert len(m.plan_buckets(spans, cap_elems=10**9)) == 1


def test_cabi_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "mingpt_b200.h")).read()
    declared = set(re.findall(r"\b(mgb_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(handle, name), name
    assert declared == set(_lib.SIGNATURES) | set(_lib.INTROSPECTION)
    assert handle.mgb_version() == 100
    assert ctypes.sizeof(_lib.GemmArgs) == 232 and ctypes.sizeof(_lib.AdamWChunk) == 16


def test_device_window_loader_follows_the_distributed_sampler_order():
    """DeviceWindowLoader (host side only here): window starts == the indices torch's DistributedSampler hands the
    reference's DataLoader (mingpt/trainer.py:73-81), per rank, through random_split Subsets (mingpt/train.py:20-22),
    same order every epoch unless set_epoch is called (the reference never calls it)"""
    from torch.utils.data import random_split
    from torch.utils.data.distributed import DistributedSampler
    data = bytes((97 + (i * 7 + i // 5) % 26) for i in range(3000))
    ds = m.CharDataset(m.DataConfig(path=None, block_size=16, train_split=0.9, truncate=1.0), data=data)
    for rank in (0, 1, 2):
        sampler = DistributedSampler(ds, num_replicas=3, rank=rank)
        loader = m.DeviceWindowLoader(ds, 64, "cpu", sampler)
        starts = loader.window_starts()
        assert starts.tolist() == list(iter(DistributedSampler(ds, num_replicas=3, rank=rank)))
        assert len(loader) == (len(sampler) + 63) // 64
        again = loader.window_starts()
        assert torch.equal(starts, again)                       # no set_epoch -> same order (reference behaviour)
        sampler.set_epoch(1)
        assert not torch.equal(loader.window_starts(), starts)  # the fix behind GPTTrainerConfig.shuffle_each_epoch
    train, test = random_split(ds, [2000, len(ds) - 2000], generator=torch.Generator().manual_seed(3))
    sampler = DistributedSampler(test, num_replicas=1, rank=0, shuffle=False)
    loader = m.DeviceWindowLoader(test, 32, "cpu", sampler)
    starts = loader.window_starts()
    assert starts.tolist() == [test.indices[i] for i in range(len(test))]
    x0, y0 = test[5]
    s = int(starts[5])
    assert torch.equal(ds.codes[s:s + 16], x0) and torch.equal(ds.codes[s + 1:s + 17], y0)
    with pytest.raises(TypeError):
        m.DeviceWindowLoader(torch.utils.data.TensorDataset(torch.zeros(4)), 2, "cpu", None)


def test_dropout_keep_word_buffer_size_is_a_host_side_computation():
    """mgb_attention_dropout_bits_words needs no GPU: B*H*ceil(T/32) * (roundup(T,16) + roundup(T,32)) 32-bit words
    (query-major words for the forward + the bit-transposed copy for the backward, include/mingpt_b200.h)"""
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    handle = ctypes.CDLL(_lib.LIB_PATH)
    fn = handle.mgb_attention_dropout_bits_words
    fn.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
    fn.restype = ctypes.c_int
    for B, T, H in [(32, 1024, 12), (2, 136, 3), (1, 1, 1), (3, 33, 5)]:
        n = ctypes.c_int64(-1)
        assert fn(B, T, H, ctypes.addressof(n)) == 0
        kw = (T + 31) // 32
        assert n.value == B * H * kw * ((T + 15) // 16 * 16 + kw * 32)
    assert fn(0, 16, 1, ctypes.addressof(n)) != 0 and fn(1, 16, 1, None) != 0


def test_flops_per_token_matches_the_baseline_table():
    """F_tok = 6 (12 L C^2 + V C) + 6 L C T: the numbers MFU is computed from"""
    from mingpt_distributed_b200.flops import flops_per_token, mfu
    assert flops_per_token(12, 768, 50257, 1024) == 797_815_296
    assert flops_per_token(24, 1024, 50257, 1024) == 2_271_713_280
    assert flops_per_token(48, 1600, 50257, 1024) == 9_801_686_400
    assert flops_per_token(6, 192, 65, 256) == 17_769_600
    assert flops_per_token(3, 48, 65, 128) == 626_976
    # the as-written graph: two output projections, full-square attention
    assert flops_per_token(12, 768, 50257, 1024, "reference") == 6 * (13 * 12 * 768 ** 2 + 50257 * 768) + 12 * 12 * 768 * 1024
    assert abs(mfu(2_820_202, 797_815_296) - 1.0) < 1e-6        # tokens/s at 100 % of 2.25 PFLOP/s


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` is CPU-only (the oracle port of the reference's as-written graph): it must run
    without a GPU and print one JSON line with the contract's keys"""
    import json
    import subprocess
    import sys
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "gpt-nano",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "train_tokens_per_s" and line["unit"] == "tokens/s"
    assert line["value"] > 0 and line["higher_is_better"] is True and line["n_gpus"] == 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] in ("port", "reference") and line["cpu_baseline"]["cores"] >= 1
