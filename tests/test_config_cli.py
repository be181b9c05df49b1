"""Config loaders and CLI surface: CPU only."""
import json
import os

import pytest

from b200bench import cli
from b200bench.config import StrategyConfig, get_model_config, load_deepspeed_config, load_fsdp_config
from b200bench.parallel.torch_engine import effective_grad_accum

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REQUIRED = ["--strategy", "ddp", "--world-size", "1", "--rank", "0", "--tier", "A", "--seq-len", "2048",
            "--steps", "10", "--per-device-batch", "1", "--grad-accum", "4", "--results-dir", "/tmp/r"]


def test_tier_table_matches_reference():
    # TH:157-179: A = 32000/1024/16/16, B = 32000/2048/32/32, block_size = seq_len
    a, b = get_model_config("A", 2048), get_model_config("B", 4096)
    assert (a["vocab_size"], a["n_embd"], a["n_head"], a["n_layer"], a["block_size"]) == (32000, 1024, 16, 16, 2048)
    assert (b["vocab_size"], b["n_embd"], b["n_head"], b["n_layer"], b["block_size"]) == (32000, 2048, 32, 32, 4096)
    with pytest.raises((ValueError, KeyError)):
        get_model_config("Z", 128)


@pytest.mark.parametrize("stage", ["zero2", "zero3"])
def test_deepspeed_json_batch_keys_and_optimizer(stage):
    path = os.path.join(ROOT, "configs", "deepspeed", f"{stage}.json")
    raw = json.load(open(path))
    cfg = load_deepspeed_config(path, micro_batch=3, grad_accum=4, world_size=8)
    # the three batch keys are replaced by integers computed from the CLI (TH:251-262)
    assert cfg.raw["train_micro_batch_size_per_gpu"] == 3
    assert cfg.raw["gradient_accumulation_steps"] == 4
    assert cfg.raw["train_batch_size"] == 3 * 4 * 8
    assert all(isinstance(cfg.raw[k], int) for k in ("train_batch_size", "train_micro_batch_size_per_gpu",
                                                     "gradient_accumulation_steps"))
    opt = raw["optimizer"]["params"]
    assert cfg.optim.weight_decay == pytest.approx(float(opt["weight_decay"]))
    assert cfg.optim.grad_clip == pytest.approx(float(raw["gradient_clipping"]))
    assert cfg.optim.warmup_steps == int(raw["scheduler"]["params"]["warmup_num_steps"])
    assert cfg.optim.lr == pytest.approx(float(raw["scheduler"]["params"]["warmup_max_lr"]))
    assert cfg.reduce_bucket_size == int(float(raw["zero_optimization"]["reduce_bucket_size"]))
    if stage == "zero3":
        assert cfg.stage3_max_live_parameters == int(float(raw["zero_optimization"]["stage3_max_live_parameters"]))


def test_deepspeed_offload_is_rejected(tmp_path):
    raw = json.load(open(os.path.join(ROOT, "configs", "deepspeed", "zero3.json")))
    raw["zero_optimization"]["offload_param"] = {"device": "cpu"}
    p = tmp_path / "z.json"
    p.write_text(json.dumps(raw))
    with pytest.raises(ValueError, match="offload"):
        load_deepspeed_config(str(p), 1, 1, 2)


def test_fsdp_yaml_is_applied(tmp_path):
    cfg = load_fsdp_config(os.path.join(ROOT, "configs", "fsdp", "fsdp_config.yaml"))
    assert cfg.fsdp_sharding_strategy == "full_shard"
    assert cfg.fsdp_auto_wrap_policy == "transformer_block"
    assert cfg.fsdp_forward_prefetch is True
    assert cfg.fsdp_reshard_after_forward == "auto"
    p = tmp_path / "f.yaml"
    p.write_text("fsdp_config:\n  reshard_after_forward: true\n  forward_prefetch: false\n")
    cfg = load_fsdp_config(str(p))
    assert cfg.fsdp_reshard_after_forward is True and cfg.fsdp_forward_prefetch is False
    p.write_text("fsdp_config:\n  reshard_after_forward: sometimes\n")
    with pytest.raises(ValueError, match="reshard_after_forward"):
        load_fsdp_config(str(p))
    p.write_text("fsdp_config:\n  cpu_offload: true\n")
    with pytest.raises(ValueError, match="cpu_offload"):
        load_fsdp_config(str(p))
    assert isinstance(load_fsdp_config(""), StrategyConfig)      # no file: defaults


def test_grad_accum_semantics():
    # only the ZeRO strategies honour --grad-accum in reference mode
    assert effective_grad_accum("ddp", 4, "reference") == 1
    assert effective_grad_accum("fsdp", 4, "reference") == 1
    assert effective_grad_accum("zero2", 4, "reference") == 4
    assert effective_grad_accum("zero3", 4, "reference") == 4
    assert effective_grad_accum("ddp", 4, "uniform") == 4


def test_cli_reference_flags_and_defaults():
    a = cli.parse_args(REQUIRED + ["--synthetic", "--fsdp-config", "x.yaml"])
    # defaults of the reference parser (TH:465-498)
    assert (a.local_rank, a.master_addr, a.master_port, a.warmup_steps) == (0, "localhost", 29500, 5)
    assert a.synthetic and a.fsdp_config == "x.yaml" and a.deepspeed_config is None
    # extensions keep reference behaviour by default
    assert (a.attn, a.dropout, a.grad_accum_mode, a.timer, a.comm) == ("bidir", 0.1, "reference", "device", "fused")


def test_cli_rejects_zero_without_config_and_unknown_strategy():
    argv = [x if x != "ddp" else "zero2" for x in REQUIRED]
    with pytest.raises(ValueError, match="deepspeed-config"):          # TH:501-502
        cli.parse_args(argv)
    assert cli.parse_args(argv + ["--deepspeed-config", "c.json"]).deepspeed_config == "c.json"
    with pytest.raises(SystemExit):
        cli.parse_args([x if x != "ddp" else "pipeline" for x in REQUIRED])
    with pytest.raises(SystemExit):                                    # a required flag is missing
        cli.parse_args(REQUIRED[:-2])


def test_engine_knobs_come_from_the_strategy_files(tmp_path):
    z3 = load_deepspeed_config(os.path.join(ROOT, "configs", "deepspeed", "zero3.json"), 1, 4, 8)
    assert z3.param_gather == "stream" and z3.gather_ctas == 0 and z3.comm_ctas == 0
    raw = json.load(open(os.path.join(ROOT, "configs", "deepspeed", "zero3.json")))
    raw["b200bench"].update({"param_gather": "gemm", "gather_ctas": 64, "comm_ctas": 16})
    p = tmp_path / "z3.json"
    p.write_text(json.dumps(raw))
    cfg = load_deepspeed_config(str(p), 1, 4, 8)
    assert (cfg.param_gather, cfg.gather_ctas, cfg.comm_ctas) == ("gemm", 64, 16)
    raw["b200bench"]["param_gather"] = "ring"
    p.write_text(json.dumps(raw))
    with pytest.raises(ValueError, match="param_gather"):
        load_deepspeed_config(str(p), 1, 4, 8)
    raw["b200bench"].update({"param_gather": "stream", "gather_ctas": 100000})
    p.write_text(json.dumps(raw))
    with pytest.raises(ValueError, match="gather_ctas"):
        load_deepspeed_config(str(p), 1, 4, 8)
    f = load_fsdp_config(os.path.join(ROOT, "configs", "fsdp", "fsdp_config.yaml"))
    assert f.param_gather == "stream" and f.gather_ctas == 0
    assert StrategyConfig().param_gather is None                      # plain ddp: engine default


# ---- round 2: every key of a strategy file is applied, pinned to its one supported value, or rejected ----------
def test_strategy_files_reject_what_would_be_ignored(tmp_path):
    import json

    import yaml

    from b200bench.config import load_deepspeed_config, load_fsdp_config

    ref = "/root/reference/configs"
    if os.path.isdir(ref):                      # the reference's own files load unchanged
        for n in ("zero2", "zero3"):
            load_deepspeed_config(f"{ref}/deepspeed/{n}.json", 1, 4, 8)
        f = load_fsdp_config(f"{ref}/fsdp/fsdp_config.yaml")
        assert f.fsdp_sharding_strategy == "full_shard" and len(f.notes) == 2      # size_based, mixed_precision: false
    raw = json.load(open(os.path.join(ROOT, "configs", "deepspeed", "zero3.json")))

    def ds(mut):
        r = json.loads(json.dumps(raw))
        mut(r)
        p = tmp_path / "z.json"
        p.write_text(json.dumps(r))
        return load_deepspeed_config(str(p), 1, 4, 8)

    with pytest.raises(ValueError, match="contiguous_gradients"):
        ds(lambda r: r["zero_optimization"].update(contiguous_gradients=False))
    with pytest.raises(ValueError, match="not supported"):
        ds(lambda r: r["zero_optimization"].update(zero_hpz_partition_size=2))
    with pytest.raises(ValueError, match="not supported"):
        ds(lambda r: r.update(activation_checkpointing={"partition_activations": True}))
    with pytest.raises(ValueError, match="bf16"):
        ds(lambda r: r["bf16"].update(enabled=False))
    with pytest.raises(ValueError, match="stage"):
        ds(lambda r: r["zero_optimization"].update(stage=1))
    c = ds(lambda r: r["zero_optimization"].update(overlap_comm=False, stage3_prefetch_bucket_size=2.6e7))
    assert c.overlap_comm is False and c.stage3_prefetch_bucket_size == 26000000

    def fs(**kw):
        p = tmp_path / "f.yaml"
        p.write_text(yaml.safe_dump({"fsdp_config": kw}))
        return load_fsdp_config(str(p))

    assert fs(sharding_strategy="SHARD_GRAD_OP").fsdp_sharding_strategy == "shard_grad_op"
    assert fs(backward_prefetch="BACKWARD_POST", forward_prefetch=False).fsdp_forward_prefetch is False
    for bad in (dict(sharding_strategy="hybrid_shard"), dict(backward_prefetch="sometimes"), dict(use_orig_params=True),
                dict(mixed_precision="fp16"), dict(auto_wrap_policy="by_name")):
        with pytest.raises(ValueError):
            fs(**bad)


def test_fsdp_sharding_strategy_selects_the_communication_pattern():
    from b200bench.config import StrategyConfig
    from b200bench.parallel.torch_engine import comm_strategy

    c = StrategyConfig()
    assert comm_strategy("fsdp", c) == "fsdp" and comm_strategy("zero3", c) == "zero3"
    c.fsdp_sharding_strategy = "shard_grad_op"
    assert comm_strategy("fsdp", c) == "zero2" and comm_strategy("ddp", c) == "ddp"
    c.fsdp_sharding_strategy = "no_shard"
    assert comm_strategy("fsdp", c) == "ddp"
