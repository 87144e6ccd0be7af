"""Distributed plumbing on CPU / gloo, world_size 2 (BASELINE.json config #1):
sharded (ZeRO-3, resident and streamed) = ZeRO-2 = replicated (DDP) = single process on the same global
batch; checkpoint save -> load round trip; GPT-2 small-shaped DPO runs end to end through the CLI."""
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dpo_worker(rank, world, port, sharding, resident, steps, accum, out):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    torch.set_num_threads(2)
    from dla_b200.algo import dpo_step
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models import CausalLM, build_units, resolve_model_config
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.engine import Engine
    from dla_b200.parallel.pg import init_distributed, shutdown

    ctx = init_distributed(device="cpu")
    cfg = resolve_model_config("tiny-gpt2")
    units = build_units(cfg)
    pol = Engine(cfg, units, ctx, sharding=sharding, trainable=True, checkpoint=True, dtype=torch.float32, seed=7, resident=resident)
    ref = Engine(cfg, units, ctx, sharding=sharding, trainable=False, checkpoint=False, dtype=torch.float32, seed=7, resident=resident)
    mp_, mr = CausalLM(cfg, pol), CausalLM(cfg, ref)
    opt = AdamWConfig(lr=5e-3, weight_decay=0.01, max_grad_norm=1.0, eps=1e-6)
    G = 4                                             # global pairs per micro-step
    losses = []
    for step in range(steps):
        for a in range(accum):
            full = synthetic_preference_batch(G, 12, cfg.vocab_size, seed=100 + step * accum + a, ragged=True)
            lo, hi = rank * G // world, (rank + 1) * G // world
            batch = {s: {k: v[lo:hi] for k, v in full[s].items()} for s in full}
            pol.sync_grads = (a == accum - 1)
            loss, _ = dpo_step(mp_, mr, batch, beta=0.5)
            loss.backward()
            pol.finish_backward()
            losses.append(float(ctx.all_reduce(loss.detach().clone(), "mean")))
        pol.optimizer_step(opt, grad_divisor=float(world * accum))
    flat = torch.cat([pol.gathered_unit(u.name)[p.name].reshape(-1) for u in units for p in u.params])
    if rank == 0:
        torch.save({"losses": losses, "params": flat.clone(), "gnorm": float(pol.last_grad_norm)}, out)
    shutdown()


def _same(a, b, tol=2e-3):
    """Adam divides by sqrt(v): elements whose gradient is ~0 amplify summation-order noise, so compare in
    norm (and bound the outliers) instead of element-wise."""
    rel = float((a - b).norm() / b.norm())
    frac = float(((a - b).abs() > 1e-4).float().mean())
    return rel < tol and frac < 5e-3


def _run(tmp_path, world, sharding, resident=None, steps=3, accum=1, tag=""):
    out = tmp_path / f"res_{sharding}_{world}_{resident}_{accum}{tag}.pt"
    port = _free_port()
    if world == 1:
        p = mp.get_context("spawn").Process(target=_dpo_worker, args=(0, 1, port, sharding, resident, steps, accum, str(out)))
        p.start()
        p.join(300)
        assert p.exitcode == 0
    else:
        mp.spawn(_dpo_worker, args=(world, port, sharding, resident, steps, accum, str(out)), nprocs=world, join=True)
    return torch.load(out)


@pytest.mark.timeout(600)
def test_sharded_replicated_single_agree(tmp_path):
    single = _run(tmp_path, 1, "ddp")
    ddp = _run(tmp_path, 2, "ddp")
    z2 = _run(tmp_path, 2, "zero2")
    z3 = _run(tmp_path, 2, "zero3", resident=True)
    z3s = _run(tmp_path, 2, "zero3", resident=False)
    assert single["losses"][0] == pytest.approx(0.693147, abs=1e-5)     # policy == reference at step 0
    assert single["losses"][1] != pytest.approx(single["losses"][0], abs=1e-6)   # the update changed the policy
    for name, res in (("ddp", ddp), ("zero2", z2), ("zero3", z3), ("zero3-streamed", z3s)):
        assert res["losses"] == pytest.approx(single["losses"], abs=2e-5), name
        assert _same(res["params"], single["params"]), name
        assert res["gnorm"] == pytest.approx(single["gnorm"], rel=1e-4), name


@pytest.mark.timeout(600)
def test_gradient_accumulation_matches_large_batch(tmp_path):
    a = _run(tmp_path, 2, "zero3", resident=True, steps=2, accum=2, tag="acc")
    b = _run(tmp_path, 2, "zero3", resident=False, steps=2, accum=2, tag="acc")
    c = _run(tmp_path, 1, "ddp", steps=2, accum=2, tag="acc")
    assert _same(a["params"], c["params"])
    assert _same(b["params"], c["params"])


def _ckpt_worker(rank, world, port, root):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    torch.set_num_threads(2)
    from dla_b200.ckpt import CheckpointManager, load_model_dir
    from dla_b200.data import ByteTokenizer
    from dla_b200.models import build_units, resolve_model_config
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.engine import Engine
    from dla_b200.parallel.pg import init_distributed, shutdown

    ctx = init_distributed(device="cpu")
    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    eng = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=3, resident=False, name="policy")
    for u in units:       # fake one optimizer step so the moments are non-trivial
        eng.gshard[u.name].normal_(generator=torch.Generator().manual_seed(rank))
    eng.optimizer_step(AdamWConfig(lr=1e-2), grad_divisor=1.0)
    mgr = CheckpointManager(root, ctx, keep_last_n=2)
    for step in (1, 2, 3):
        mgr.save(f"step_{step}", {"policy": eng}, {"global_step": step}, tokenizer=ByteTokenizer(512))
    mgr.save("final", {"policy": eng}, {"global_step": 3}, tokenizer=ByteTokenizer(512))
    # exact resume into a fresh engine
    eng2 = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=99, resident=False, name="policy")
    state = mgr.load(mgr.resolve("latest"), {"policy": eng2})
    assert state["global_step"] == 3 and eng2.step_count == 1
    for u in units:
        assert torch.equal(eng2.master[u.name], eng.master[u.name])
        assert torch.equal(eng2.m[u.name], eng.m[u.name]) and torch.equal(eng2.v[u.name], eng.v[u.name])
    # weights-only load through the HF-named safetensors into a REPLICATED engine (different layout)
    eng3 = Engine(cfg, units, ctx, sharding="ddp", dtype=torch.float32, seed=5, trainable=False, name="other")
    assert load_model_dir(eng3, Path(root) / "latest") > 0
    for u in units:
        a, b = eng.gathered_unit(u.name), eng3.gathered_unit(u.name)
        for k in a:
            assert torch.equal(a[k], b[k]), (u.name, k)
    shutdown()


def _async_ckpt_worker(rank, world, port, root):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    torch.set_num_threads(2)
    from dla_b200.ckpt import CheckpointManager
    from dla_b200.data import ByteTokenizer
    from dla_b200.models import build_units, resolve_model_config
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.engine import Engine
    from dla_b200.parallel.pg import init_distributed, shutdown

    ctx = init_distributed(device="cpu")
    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    eng = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=3, resident=False, name="policy")
    mgr = CheckpointManager(root, ctx, keep_last_n=2, async_write=True)
    snaps = {}
    for step in (1, 2, 3):
        for u in units:
            eng.gshard[u.name].normal_(generator=torch.Generator().manual_seed(100 * step + rank))
        eng.optimizer_step(AdamWConfig(lr=1e-2), grad_divisor=1.0)
        snaps[step] = {u.name: eng.master[u.name].clone() for u in units}
        mgr.save(f"step_{step}", {"policy": eng}, {"global_step": step}, tokenizer=ByteTokenizer(512))
        # training goes on while the writer thread works: the snapshot must not see these later updates
    mgr.wait()
    ctx.barrier()
    eng2 = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=99, resident=False, name="policy")
    state = mgr.load(mgr.resolve("latest"), {"policy": eng2})
    assert state["global_step"] == 3 and eng2.step_count == 3
    for u in units:
        assert torch.equal(eng2.master[u.name], snaps[3][u.name])
    eng3 = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=98, resident=False, name="policy")
    mgr.load(Path(root) / "step_2", {"policy": eng3})
    for u in units:
        assert torch.equal(eng3.master[u.name], snaps[2][u.name])
    shutdown()


@pytest.mark.timeout(600)
def test_async_checkpoint_snapshots_then_writes_in_the_background(tmp_path):
    """``CheckpointManager(async_write=True)``: ``save`` returns after the snapshot, files appear from the writer thread,
    each checkpoint holds the state of ITS step although training continued, and the directory is complete after wait()."""
    root = tmp_path / "ckpt"
    mp.spawn(_async_ckpt_worker, args=(2, _free_port(), str(root)), nprocs=2, join=True)
    names = sorted(p.name for p in root.iterdir())
    assert names == ["latest", "step_2", "step_3"], names
    assert os.readlink(root / "latest") == "step_3"
    assert not list((root / "step_3" / "shards").glob(".done_*"))


def _resume_other_world_worker(rank, world, port, root):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    torch.set_num_threads(2)
    from dla_b200.ckpt import CheckpointManager
    from dla_b200.models import build_units, resolve_model_config
    from dla_b200.parallel.engine import Engine
    from dla_b200.parallel.pg import init_distributed, shutdown

    ctx = init_distributed(device="cpu")
    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    eng = Engine(cfg, units, ctx, sharding="zero3", dtype=torch.float32, seed=1234, resident=False, name="policy")
    mgr = CheckpointManager(root, ctx)
    state = mgr.load(mgr.resolve("latest"), {"policy": eng})
    assert state["global_step"] == 3
    full = {u.name: {k: v.clone() for k, v in eng.gathered_unit(u.name).items()} for u in units}
    if rank == 0:
        torch.save(full, Path(root) / f"loaded_world{world}.pt")
    shutdown()


@pytest.mark.timeout(900)
def test_checkpoint_written_at_world_2_loads_at_world_1_and_3(tmp_path):
    """A step-N checkpoint is loadable at a DIFFERENT world size: the shard files no longer match, so the weights come
    from the consolidated HF-named safetensors (optimizer state restarts)."""
    root = tmp_path / "ckpt"
    mp.spawn(_ckpt_worker, args=(2, _free_port(), str(root)), nprocs=2, join=True)
    for world in (1, 3):
        mp.spawn(_resume_other_world_worker, args=(world, _free_port(), str(root)), nprocs=world, join=True)
    a, b = torch.load(root / "loaded_world1.pt"), torch.load(root / "loaded_world3.pt")
    from safetensors.torch import load_file

    ref = load_file(str(root / "final" / "model.safetensors"))
    n = 0
    for unit in a:
        for k in a[unit]:
            assert torch.equal(a[unit][k], b[unit][k]), (unit, k)
            n += a[unit][k].numel()
    assert n == sum(t.numel() for t in ref.values())


@pytest.mark.timeout(600)
def test_checkpoint_layout_roundtrip_and_keep_last_n(tmp_path):
    root = tmp_path / "ckpt"
    mp.spawn(_ckpt_worker, args=(2, _free_port(), str(root)), nprocs=2, join=True)
    names = sorted(p.name for p in root.iterdir())
    assert names == ["final", "latest", "step_2", "step_3"], names          # keep_last_n=2 pruned step_1
    assert os.readlink(root / "latest") == "final"
    for f in ("config.json", "generation_config.json", "model.safetensors", "trainer_state.json", "dla_tokenizer.json",
              "shards/meta.json", "shards/rank_0.policy.safetensors", "shards/rank_1.policy.safetensors"):
        assert (root / "final" / f).exists(), f
    from dla_b200.models import resolve_model_config

    assert resolve_model_config(str(root / "latest")).hidden_size == 256   # loadable as a model directory


@pytest.mark.timeout(900)
def test_gpt2_small_dpo_cli_gloo_world2(tmp_path):
    """The shipped plumbing config end to end: launcher -> 2 gloo ranks -> ZeRO-3 GPT-2 small DPO."""
    cfg = ROOT / "config" / "dpo_gpt2_cpu.yaml"
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "dla_b200.cli.launch", "--cpu", "--nproc", "2", "--port", str(_free_port()), "-m",
           "dla_b200.cli.train_dpo", "--config", str(cfg), "--override", f"logging.output_dir={tmp_path}/ck",
           "--override", f"logging.log_dir={tmp_path}/logs", "--override", "optimization.max_train_steps=2",
           "--override", "model.max_seq_length=32"]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=850)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "DPO training complete" in res.stdout
    lines = [json.loads(l) for l in (tmp_path / "logs" / "metrics.jsonl").read_text().splitlines()]
    assert any("train/loss" in l for l in lines)
    assert (tmp_path / "ck" / "latest" / "model.safetensors").exists()


@pytest.mark.timeout(120)
def test_launcher_takes_ranks_down_and_restarts_from_latest(tmp_path):
    """a rank that dies must not leave its peers hanging, and ``--max-restarts`` relaunches the job with
    ``--resume latest`` on a fresh rendezvous port."""
    import subprocess
    import time

    # the stand-in training script is importable under a name of its own (a `tests` package elsewhere on the path would
    # shadow `-m tests.launch_crash_helper`)
    (tmp_path / "dla_launch_crash_helper.py").write_text((ROOT / "tests" / "launch_crash_helper.py").read_text())
    env = dict(os.environ, PYTHONPATH=f"{ROOT}{os.pathsep}{tmp_path}", CUDA_VISIBLE_DEVICES="")
    marker = tmp_path / "marker"
    t0 = time.monotonic()
    res = subprocess.run([sys.executable, "-m", "dla_b200.cli.launch", "--nproc", "2", "--cpu", "--port", str(_free_port()),
                          "--max-restarts", "1", "-m", "dla_launch_crash_helper", "--marker", str(marker)],
                         cwd=ROOT, env=env, capture_output=True, text=True, timeout=100)
    assert res.returncode == 0, res.stdout + res.stderr
    assert time.monotonic() - t0 < 45, "the surviving rank was not taken down promptly"
    assert "restarting 2 rank(s) with --resume latest" in res.stdout
    ports = {(tmp_path / f"done_{r}").read_text() for r in range(2)}
    assert len(ports) == 1
    # without restarts the failing rank's exit code propagates
    marker.unlink()
    res = subprocess.run([sys.executable, "-m", "dla_b200.cli.launch", "--nproc", "2", "--cpu", "--port", str(_free_port()),
                          "-m", "dla_launch_crash_helper", "--marker", str(marker)],
                         cwd=ROOT, env=env, capture_output=True, text=True, timeout=100)
    assert res.returncode == 7


def _ref_meta_worker(rank, world, port):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), DLA_REF_META_INIT="1")
    import warnings

    import torch.distributed as dist

    sys.path.insert(0, str(ROOT))
    from baseline import reference_arm as ra

    if ra.locate() is None:
        return
    warnings.simplefilter("ignore")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    runner = ra.ReferenceDPO("tiny-llama", torch.device("cpu"), world, sharded=True)
    assert runner.meta_init
    assert not any(p.is_meta for m in (runner.policy, runner.ref) for p in m.parameters())
    inv = runner.policy.model.rotary_emb.inv_freq
    assert not inv.is_meta and float(inv[0]) == 1.0 and bool(torch.isfinite(inv).all())
    gains = [p for n, p in runner.policy.named_parameters() if n.endswith("norm.weight")]
    assert gains and all(bool((g.to_local() == 1).all()) for g in gains)
    vocab = ra.llama_hf_config("tiny-llama").vocab_size
    for i in range(2):
        loss = runner.step(ra.synthetic_preference_batch(2, 32, vocab, seed=i + 10 * rank))
        assert bool(torch.isfinite(loss))
    dist.destroy_process_group()


def test_reference_arm_meta_init_shards_before_materialising():
    """The 70B reference arm builds both replicas on the meta device, shards them with FSDP2 and only then
    allocates; the same code path is forced here on the tiny preset over gloo."""
    if not (ROOT / "baseline" / "_ref").exists() and not Path("/root/reference").exists():
        pytest.skip("reference tree not present")
    mp.spawn(_ref_meta_worker, args=(2, _free_port()), nprocs=2, join=True)


def test_reference_arm_param_count_formula():
    sys.path.insert(0, str(ROOT))
    from baseline import reference_arm as ra

    assert abs(ra.param_bytes(ra.llama_hf_config("llama3-8b")) / 2 - 8.03e9) < 0.01e9
    assert abs(ra.param_bytes(ra.llama_hf_config("llama3-70b")) / 2 - 70.55e9) < 0.05e9


@pytest.mark.timeout(900)
def test_clipped_ppo_gloo_world2_uneven_batches_and_global_early_stop(tmp_path):
    """Clipped PPO (value head, GAE, mini-batches) and REINFORCE on 2 gloo ranks with a batch size that does not split
    evenly (5 prompts; the loop trims the global batch to equal shares, since the cross-rank gathers carry one row count)
    and a ``target_kl`` so small that the early stop fires: the stop decision and the number of mini-batches must be the
    same on both ranks, or the next collective hangs.  The job has to finish, for both algorithms."""
    import yaml

    prompts = tmp_path / "prompts.jsonl"
    prompts.write_text("\n".join(json.dumps({"prompt": f"Tell me about {i}"}) for i in range(10)) + "\n")
    base = {"seed": 5, "model": {"policy_model_name_or_path": "tiny-llama", "reference_model_name_or_path": "tiny-llama", "max_seq_length": 32},
            "reward_model": {"base_model_name_or_path": "tiny-llama"},
            "ppo": {"steps": 2, "batch_size": 5, "learning_rate": 1e-4, "kl_coef": 0.1,
                    "generation_params": {"max_new_tokens": 4, "temperature": 0.7, "top_p": 0.9, "do_sample": True}},
            "sampling": {"source": "local", "prompt_path": str(prompts)},
            "hardware": {"device": "cpu"}}
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    for algo, extra in (("ppo", {"algo": "ppo", "mini_batch_size": 2, "epochs": 2, "target_kl": 1e-12}), ("reinforce", {})):
        cfg = json.loads(json.dumps(base))
        cfg["ppo"].update(extra)
        cfg["logging"] = {"output_dir": str(tmp_path / algo / "ck"), "log_dir": str(tmp_path / algo / "logs"), "log_every_steps": 1}
        path = tmp_path / f"rlhf_{algo}.yaml"
        path.write_text(yaml.safe_dump(cfg))
        cmd = [sys.executable, "-m", "dla_b200.cli.launch", "--cpu", "--nproc", "2", "--port", str(_free_port()), "-m",
               "dla_b200.cli.train_rlhf", "--config", str(path)]
        res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=400)
        assert res.returncode == 0, f"{algo}:\n" + res.stdout[-2500:] + res.stderr[-2500:]
        assert "RLHF PPO loop complete" in res.stdout
        lines = [json.loads(l) for l in (tmp_path / algo / "logs" / "metrics.jsonl").read_text().splitlines()]
        assert any("train/kl" in l for l in lines)


@pytest.mark.timeout(900)
def test_sft_reward_distill_cli_gloo_world2(tmp_path):
    """The other trainers on 2 gloo ranks (ZeRO-3), with their evaluation loops (cross-rank loss gathers) and checkpoints:
    SFT -> reward model from the SFT checkpoint -> distillation (ensemble KL) from it."""
    import yaml

    sft = [{"prompt": f"Question {i}: what is {i}+{i}?", "response": f"It is {2 * i}."} for i in range(22)]      # 22: not a multiple of 4
    pref = [{"prompt": f"Say {i}", "chosen": f"{i} indeed", "rejected": "no"} for i in range(22)]
    roll = [{"prompt": f"Tell me about {i}", "teacher_response": f"{i} is a number.", "reward": 0.1 * i} for i in range(10)]
    for name, rows in (("sft.jsonl", sft), ("pref.jsonl", pref), ("roll.jsonl", roll)):
        (tmp_path / name).write_text("\n".join(json.dumps(r) for r in rows) + "\n")
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")

    def log(phase):
        return {"output_dir": str(tmp_path / "ck" / phase), "log_dir": str(tmp_path / "logs" / phase), "log_every_steps": 1,
                "save_every_steps": 2, "eval_every_steps": 1, "keep_last_n": 2, "use_wandb": False}

    def run(module, cfg, done):
        path = tmp_path / f"{module.rsplit('.', 1)[-1]}.yaml"
        path.write_text(yaml.safe_dump(cfg))
        cmd = [sys.executable, "-m", "dla_b200.cli.launch", "--cpu", "--nproc", "2", "--port", str(_free_port()), "-m", module,
               "--config", str(path)]
        res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=400)
        assert res.returncode == 0, f"{module}:\n" + res.stdout[-2500:] + res.stderr[-2500:]
        assert done in res.stdout, res.stdout[-1500:]

    hw = {"device": "cpu", "sharding": "zero3"}
    run("dla_b200.cli.train_sft", {
        "seed": 1, "model": {"model_name_or_path": "tiny-llama", "max_seq_length": 48},
        "data": {"source": "local", "train_path": str(tmp_path / "sft.jsonl"), "eval_path": str(tmp_path / "sft.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "warmup_steps": 1, "max_train_steps": 3},
        "logging": log("sft"), "hardware": {**hw, "gradient_accumulation_steps": 2}}, "Training complete")
    latest = tmp_path / "ck" / "sft" / "latest"
    assert (latest / "model.safetensors").exists()
    run("dla_b200.cli.train_reward", {
        "seed": 2, "model": {"base_model_name_or_path": str(latest), "pooling": "mean", "dropout": 0.1, "max_seq_length": 48},
        "data": {"source": "local", "train_path": str(tmp_path / "pref.jsonl"), "eval_path": str(tmp_path / "pref.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "max_train_steps": 2},
        "logging": log("reward"), "hardware": hw}, "Reward model training complete")
    assert (tmp_path / "ck" / "reward" / "latest" / "pytorch_model.bin").exists()
    run("dla_b200.cli.train_distill", {
        "seed": 4, "model": {"student_model_name_or_path": "tiny-llama", "max_seq_length": 48},
        "data": {"teacher_samples_path": str(tmp_path / "roll.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "max_train_steps": 2, "temperature": 2.0},
        "distill": {"use_kl": True, "on_policy": True, "teacher_model_names_or_paths": [str(latest), "tiny-llama"]},
        "logging": log("distill"), "hardware": hw}, "Distillation complete")
    m = [json.loads(l) for l in (tmp_path / "logs" / "sft" / "metrics.jsonl").read_text().splitlines()]
    assert any("eval/loss" in x for x in m) and any("train/loss" in x for x in m)
