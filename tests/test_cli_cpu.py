"""Every phase entrypoint end to end on CPU with tiny random-init models and synthetic / local JSONL data:
SFT -> reward -> DPO (from the SFT checkpoint) -> teacher rollouts -> distillation (CE and KL) -> RLHF
(reinforce and ppo) -> eval.  Mirrors the reference's phase chain (R/README.md:40-46) and its file-system
hand-off between phases (checkpoints/<phase>/latest, JSONL rollouts)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import yaml

ROOT = Path(__file__).resolve().parents[1]


_DEVICE = ["cpu"]        # the chain below runs once on CPU and once (marked gpu) on a CUDA device


def _run(module, *args, timeout=600):
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="2")
    if _DEVICE[0] == "cpu":
        env["CUDA_VISIBLE_DEVICES"] = ""
    res = subprocess.run([sys.executable, "-m", module, *args], cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert res.returncode == 0, f"{module} failed:\n{res.stdout[-3000:]}\n{res.stderr[-3000:]}"
    return res.stdout


def _write(path: Path, obj) -> str:
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text(yaml.safe_dump(obj))
    return str(path)


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    d = tmp_path_factory.mktemp("phases")
    sft = [{"prompt": f"Question {i}: what is {i}+{i}?", "response": f"It is {2 * i}."} for i in range(24)]
    pref = [{"prompt": f"Say {i}", "chosen": f"{i} indeed", "rejected": "no"} for i in range(24)]
    prompts = [{"prompt": f"Tell me about {i}"} for i in range(8)]
    for name, rows in (("sft.jsonl", sft), ("pref.jsonl", pref), ("prompts.jsonl", prompts)):
        (d / name).write_text("\n".join(json.dumps(r) for r in rows) + "\n")
    return d


def _log(d: Path, phase: str, **kw):
    base = {"output_dir": str(d / "ckpt" / phase), "log_dir": str(d / "logs" / phase), "log_every_steps": 1, "save_every_steps": 2,
            "eval_every_steps": 2, "keep_last_n": 2, "use_wandb": False}
    base.update(kw)
    return base


@pytest.mark.timeout(600)
@pytest.mark.parametrize("device", ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)])
def test_sft_with_sequence_packing(work, device):
    """``data.packing: true`` (R/config/sft_config.yaml:16) packs several documents per row; on CUDA the flash kernels
    take the document boundaries, so the flag no longer falls back to padded rows there."""
    _DEVICE[0] = device
    d = work / f"pack_{device}"
    d.mkdir(exist_ok=True)
    cfg = _write(d / "sft_pack.yaml", {
        "seed": 1, "model": {"model_name_or_path": "tiny-llama", "max_seq_length": 160},
        "data": {"source": "local", "train_path": str(work / "sft.jsonl"), "num_workers": 0, "packing": True},
        "optimization": {"micro_batch_size": 6, "learning_rate": 1e-3, "warmup_steps": 1, "max_train_steps": 3},
        "logging": _log(d, "sft_pack", save_every_steps=0), "hardware": {"device": device}})
    out = _run("dla_b200.cli.train_sft", "--config", cfg, "--override", "logging.torch_profiler={start_step: 1, steps: 1}")
    assert "Training complete" in out and "torch.profiler window" in out
    prof_dir = d / "logs" / "sft_pack" / "profiler"
    assert (prof_dir / "trace_rank0_step1.json").exists() and (prof_dir / "kernels_step1.txt").exists()
    metrics = [json.loads(l) for l in (d / "logs" / "sft_pack" / "metrics.jsonl").read_text().splitlines()]
    losses = [m["train/loss"] for m in metrics if "train/loss" in m]
    assert losses and all(l == l and l < 20 for l in losses)


@pytest.mark.timeout(1500)
@pytest.mark.parametrize("device", ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)])
def test_phase_chain(work, device):
    """device="cuda": the same chain through the native kernels (bf16), one process on one GPU."""
    _DEVICE[0] = device
    d = work / device
    d.mkdir(exist_ok=True)
    for f in ("sft.jsonl", "pref.jsonl", "prompts.jsonl"):
        (d / f).write_text((work / f).read_text())
    hw = {"device": device, "gradient_accumulation_steps": 2}
    # ---- SFT (cosine schedule, eval split, prompt masking)
    sft_cfg = _write(d / "sft.yaml", {
        "seed": 1, "model": {"model_name_or_path": "tiny-llama", "max_seq_length": 64},
        "data": {"source": "local", "train_path": str(d / "sft.jsonl"), "eval_path": str(d / "sft.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "warmup_steps": 1, "max_train_steps": 4, "weight_decay": 0.01},
        "logging": _log(d, "sft"), "hardware": hw})
    out = _run("dla_b200.cli.train_sft", "--config", sft_cfg)
    assert "Training complete" in out and "Trainable parameters" in out
    latest = d / "ckpt" / "sft" / "latest"
    assert (latest / "model.safetensors").exists() and (latest / "config.json").exists()
    metrics = [json.loads(l) for l in (d / "logs" / "sft" / "metrics.jsonl").read_text().splitlines()]
    assert any("eval/loss" in m for m in metrics) and any("train/loss" in m for m in metrics)
    lrs = [m["train/lr"] for m in metrics if "train/lr" in m]
    assert lrs[0] != lrs[-1]                                     # the cosine schedule moved the LR
    # ---- reward model on the SFT backbone, writes pytorch_model.bin for the PPO phase
    rw_cfg = _write(d / "reward.yaml", {
        "seed": 2, "model": {"base_model_name_or_path": str(latest), "pooling": "last_token", "dropout": 0.1, "max_seq_length": 48},
        "data": {"source": "local", "train_path": str(d / "pref.jsonl"), "eval_path": str(d / "pref.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "max_train_steps": 2},
        "logging": _log(d, "reward"), "hardware": {"device": device}})
    out = _run("dla_b200.cli.train_reward", "--config", rw_cfg)
    assert "Reward model training complete" in out
    assert (d / "ckpt" / "reward" / "latest" / "pytorch_model.bin").exists()
    # ---- DPO from the SFT checkpoint, with an ablation overlay and an override, then resume
    dpo_cfg = _write(d / "dpo.yaml", {
        "seed": 3, "model": {"policy_model_name_or_path": str(latest), "reference_model_name_or_path": str(latest), "beta": 0.1,
                             "max_seq_length": 48},
        "data": {"source": "local", "preference_path": str(d / "pref.jsonl"), "num_workers": 0},
        "optimization": {"micro_batch_size": 2, "learning_rate": 1e-4, "max_train_steps": 2},
        "logging": _log(d, "dpo"), "hardware": {"device": device}})
    out = _run("dla_b200.cli.train_dpo", "--config", dpo_cfg, "--overlay", str(ROOT / "config/ablations/high_beta_dpo.yaml"),
               "--override", "optimization.max_train_steps=2")
    assert "DPO training complete" in out
    out = _run("dla_b200.cli.train_dpo", "--config", dpo_cfg, "--resume", "latest", "--override", "optimization.max_train_steps=3")
    assert "Resumed from" in out and "step 3/3" in out
    # ---- teacher rollouts with reward scores -> JSONL
    roll = d / "rollouts.jsonl"
    _run("dla_b200.cli.generate_teacher_data", "--teacher", str(d / "ckpt" / "dpo" / "latest"), "--prompts", str(d / "prompts.jsonl"),
         "--output", str(roll), "--reward_model", str(d / "ckpt" / "reward" / "latest"), "--batch_size", "4", "--max_new_tokens", "6")
    rows = [json.loads(l) for l in roll.read_text().splitlines()]
    assert len(rows) == 8 and {"prompt", "teacher_response", "reward"} <= set(rows[0])
    # ---- distillation: CE mode, then ensemble-KL mode with two teachers
    base = {"seed": 4, "model": {"student_model_name_or_path": "tiny-llama", "max_seq_length": 48},
            "data": {"teacher_samples_path": str(roll), "num_workers": 0},
            "optimization": {"micro_batch_size": 2, "learning_rate": 1e-3, "max_train_steps": 2, "temperature": 2.0},
            "logging": _log(d, "distill"), "hardware": {"device": device}}
    out = _run("dla_b200.cli.train_distill", "--config", _write(d / "distill_ce.yaml", {**base, "distill": {"use_kl": False}}))
    assert "Distillation complete" in out
    kl = {**base, "distill": {"use_kl": True, "on_policy": True,
                              "teacher_model_names_or_paths": [str(d / "ckpt" / "dpo" / "latest"), str(latest)]}}
    out = _run("dla_b200.cli.train_distill", "--config", _write(d / "distill_kl.yaml", kl))
    assert "Distillation complete" in out
    # ---- RLHF: the reference's REINFORCE update, then clipped PPO with a value head
    rl = {"seed": 5, "model": {"policy_model_name_or_path": str(latest), "reference_model_name_or_path": str(latest), "max_seq_length": 32},
          "reward_model": {"path": str(d / "ckpt" / "reward" / "latest"), "base_model_name_or_path": str(latest)},
          "ppo": {"steps": 2, "batch_size": 4, "learning_rate": 1e-4, "kl_coef": 0.1,
                  "generation_params": {"max_new_tokens": 5, "temperature": 0.7, "top_p": 0.9, "do_sample": True}},
          "sampling": {"source": "local", "prompt_path": str(d / "prompts.jsonl")},
          "logging": {"output_dir": str(d / "ckpt" / "rlhf"), "log_dir": str(d / "logs" / "rlhf"), "log_every_steps": 1},
          "hardware": {"device": device}}
    out = _run("dla_b200.cli.train_rlhf", "--config", _write(d / "rlhf.yaml", rl))
    assert "RLHF PPO loop complete" in out
    m = [json.loads(l) for l in (d / "logs" / "rlhf" / "metrics.jsonl").read_text().splitlines()]
    assert any("train/kl" in x for x in m)
    rl["ppo"].update({"algo": "ppo", "mini_batch_size": 2, "epochs": 2, "target_kl": 6.0})
    rl["logging"]["output_dir"] = str(d / "ckpt" / "rlhf_ppo")
    out = _run("dla_b200.cli.train_rlhf", "--config", _write(d / "rlhf_ppo.yaml", rl))
    assert "RLHF PPO loop complete" in out
    # ---- eval: alignment proxies + latency grid
    ev = {"seed": 0, "models": {"sft": str(latest), "dpo": str(d / "ckpt" / "dpo" / "latest")},
          "benchmarks": {"local": {"type": "local", "prompts_path": str(d / "prompts.jsonl"), "max_samples": 4}},
          "generation": {"max_new_tokens": 4, "temperature": 0.7, "top_p": 0.9, "do_sample": True},
          "latency": {"batch_sizes": [1, 2], "seq_lengths": [16], "warmup_steps": 1, "measure_steps": 2, "decode_tokens": 2},
          "logging": {"output_path": str(d / "eval" / "results.json"), "table_path": str(d / "eval" / "summary.md")}}
    evp = _write(d / "eval.yaml", ev)
    _run("dla_b200.cli.eval_alignment", "--config", evp)
    _run("dla_b200.cli.eval_latency", "--config", evp)
    res = json.loads((d / "eval" / "results.json").read_text())
    assert set(res) == {"sft", "dpo"} and set(res["sft"]["local"]) == {"avg_length", "refusal_rate", "toxicity_proxy"}
    lat = json.loads((d / "eval" / "latency.json").read_text())
    assert len(lat["sft"]) == 2 and lat["sft"][0]["tokens_per_second"] > 0
    assert "| Model | Benchmark |" in (d / "eval" / "summary.md").read_text()
