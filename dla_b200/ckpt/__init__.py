"""Checkpoint save / resume.

The reference only saves: ``accelerator.save_state(output_dir/tag)`` every ``save_every_steps`` as
``step_N`` and once as ``final`` (``R/src/training/utils.py:99-102``, ``R/src/training/train_dpo.py:134-142``);
it cannot resume, never writes the ``latest`` dir its own downstream configs load
(``R/config/dpo_config.yaml:6-7``) nor the ``pytorch_model.bin`` the PPO script expects
(``R/src/training/train_rlhf.py:79-84``), and its README's keep-last-N does not exist.

Layout written here (superset that satisfies every consumer):

    <output_dir>/step_<N>/ | final/
        config.json, generation_config.json, tokenizer files   -> loadable as a model directory
        model.safetensors | model-0000k-of-0000n.safetensors   -> consolidated weights, HF names
        pytorch_model.bin                                       -> reward phase (backbone.* + scorer.1.*)
        shards/rank_<r>.<engine>.safetensors, shards/meta.json  -> fp32 master + Adam moments (exact resume)
        trainer_state.json                                      -> global_step, epoch, sampler, RNG, lr state
    <output_dir>/latest -> step_<N> | final                     -> symlink, switched atomically
"""
from __future__ import annotations

import json
import os
import random
import shutil
import threading
from pathlib import Path
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from dla_b200.models.io import SafetensorsReader, ShardedSafetensorsWriter, export_unit, has_pretrained_weights, import_unit
from dla_b200.parallel.engine import Engine
from dla_b200.parallel.pg import DistContext


def save_model_dir(engine: Engine, out_dir: Path, ctx: DistContext, tokenizer: Any = None,
                   reward_bin: bool = False) -> None:
    """Consolidated HF-style model directory.  Collective: every rank takes part in the unit gathers,
    rank 0 writes."""
    writer = ShardedSafetensorsWriter(out_dir) if ctx.is_main else None
    bin_state: Dict[str, torch.Tensor] = {}
    for u in engine.units:
        w = engine.gathered_unit(u.name)
        if ctx.is_main:
            tensors = export_unit(engine.cfg, u.name, w)
            writer.add(tensors)
            if reward_bin:
                for k, t in tensors.items():
                    if k.startswith("scorer."):
                        bin_state[k] = t.cpu()
                    else:
                        kk = k.split(".", 1)[1] if k.split(".", 1)[0] in ("model", "transformer") else k
                        bin_state["backbone." + kk] = t.cpu()
    if ctx.is_main:
        writer.close()
        engine.cfg.save(out_dir)
        if tokenizer is not None and hasattr(tokenizer, "save_pretrained"):
            tokenizer.save_pretrained(str(out_dir))
        if reward_bin:
            torch.save(bin_state, out_dir / "pytorch_model.bin")


def load_model_dir(engine: Engine, path: str | Path, strict: bool = False) -> int:
    """Load HF-named weights from a model directory into an engine; returns #parameters loaded."""
    reader = SafetensorsReader(path)
    if not reader.has_weights():
        return 0
    loaded = 0
    for u in engine.units:
        tensors = import_unit(engine.cfg, u.name, u, reader.get)
        if strict and len(tensors) != len(u.params):
            missing = {p.name for p in u.params} - set(tensors)
            raise KeyError(f"{path}: unit {u.name} is missing {sorted(missing)}")
        for p in u.params:
            if p.name in tensors and tuple(tensors[p.name].shape) != tuple(p.shape):
                raise ValueError(f"{path}: {u.name}.{p.name} has shape {tuple(tensors[p.name].shape)}, expected {p.shape}")
        if tensors:
            engine.load_unit(u.name, tensors)
            loaded += len(tensors)
    return loaded


def _rng_state() -> Dict[str, Any]:
    st = {"python": random.getstate(), "numpy": np.random.get_state(), "torch": torch.get_rng_state()}
    if torch.cuda.is_available():
        st["cuda"] = torch.cuda.get_rng_state()
    return st


def _set_rng_state(st: Dict[str, Any]) -> None:
    random.setstate(st["python"])
    np.random.set_state(st["numpy"])
    torch.set_rng_state(st["torch"])
    if "cuda" in st and torch.cuda.is_available():
        torch.cuda.set_rng_state(st["cuda"])


class CheckpointManager:
    def __init__(self, output_dir: str | Path, ctx: DistContext, keep_last_n: Optional[int] = None,
                 async_write: bool = False) -> None:
        self.root, self.ctx, self.keep = Path(output_dir), ctx, keep_last_n
        self.async_write = async_write
        self._thread: Optional[threading.Thread] = None
        if ctx.is_main:
            self.root.mkdir(parents=True, exist_ok=True)

    # ---------------------------------------------------------------- save
    def save(self, tag: str, engines: Dict[str, Engine], trainer_state: Dict[str, Any], tokenizer: Any = None,
             primary: Optional[str] = None, reward_bin: bool = False, save_optimizer: bool = True) -> Path:
        """``engines``: name -> Engine; ``primary`` (default: first trainable) is exported as the HF model dir."""
        self.wait()
        out = self.root / tag
        tmp = self.root / f".{tag}.tmp"
        if self.ctx.is_main:
            if tmp.exists():
                shutil.rmtree(tmp)
            (tmp / "shards").mkdir(parents=True, exist_ok=True)
        self.ctx.barrier()
        primary = primary or next((n for n, e in engines.items() if e.trainable), next(iter(engines)))
        save_model_dir(engines[primary], tmp, self.ctx, tokenizer, reward_bin=reward_bin)
        if save_optimizer:
            from safetensors.torch import save_file

            for name, eng in engines.items():
                if not eng.trainable:
                    continue
                state = {k: v.detach().cpu().contiguous() for k, v in eng.shard_state().items()}
                save_file(state, str(tmp / "shards" / f"rank_{self.ctx.rank}.{name}.safetensors"))
            torch.save(_rng_state(), tmp / "shards" / f"rng_{self.ctx.rank}.pt")
        self.ctx.barrier()
        if self.ctx.is_main:
            meta = {"world_size": self.ctx.world_size,
                    "engines": {n: {"sharding": e.sharding, "step_count": e.step_count, "trainable": e.trainable,
                                    "shard_group": e.W} for n, e in engines.items()}}
            with (tmp / "shards" / "meta.json").open("w") as fh:
                json.dump(meta, fh, indent=2)
            with (tmp / "trainer_state.json").open("w") as fh:
                json.dump(trainer_state, fh, indent=2, default=str)
            if out.exists():
                shutil.rmtree(out)
            os.replace(tmp, out)
            self._point_latest(tag)
            self._prune()
        self.ctx.barrier()
        return out

    def _point_latest(self, tag: str) -> None:
        link, tmp = self.root / "latest", self.root / ".latest.tmp"
        if tmp.is_symlink() or tmp.exists():
            tmp.unlink()
        os.symlink(tag, tmp)
        os.replace(tmp, link)

    def _prune(self) -> None:
        if not self.keep or self.keep <= 0:
            return
        steps = sorted((int(p.name.split("_")[1]), p) for p in self.root.glob("step_*") if p.name.split("_")[1].isdigit())
        for _, p in steps[:-self.keep]:
            shutil.rmtree(p, ignore_errors=True)

    def wait(self) -> None:
        if self._thread is not None:
            self._thread.join()
            self._thread = None

    # ---------------------------------------------------------------- resume
    def resolve(self, resume: Optional[str]) -> Optional[Path]:
        if not resume:
            return None
        p = self.root / "latest" if resume == "latest" else Path(resume)
        if not p.exists():
            if resume == "latest":
                return None
            raise FileNotFoundError(f"--resume {resume}: no such checkpoint")
        return p.resolve()

    def load(self, path: Path, engines: Dict[str, Engine], restore_rng: bool = True) -> Dict[str, Any]:
        """Exact resume when the shard files match this world size; otherwise weights only."""
        from safetensors.torch import load_file

        meta_p = path / "shards" / "meta.json"
        meta = json.load(meta_p.open()) if meta_p.exists() else None
        for name, eng in engines.items():
            if not eng.trainable:
                continue
            f = path / "shards" / f"rank_{self.ctx.rank}.{name}.safetensors"
            ok = meta is not None and f.exists() and meta["engines"].get(name, {}).get("shard_group") == eng.W
            if ok:
                eng.load_shard_state(load_file(str(f)), int(meta["engines"][name]["step_count"]))
            else:
                load_model_dir(eng, path)
        rng = path / "shards" / f"rng_{self.ctx.rank}.pt"
        if restore_rng and rng.exists():
            _set_rng_state(torch.load(rng, weights_only=False))
        ts = path / "trainer_state.json"
        return json.load(ts.open()) if ts.exists() else {}


def save_accelerator_state(manager: CheckpointManager, tag: str, engines: Dict[str, Engine],
                           trainer_state: Dict[str, Any], **kw: Any) -> Path:
    """Name-compatible wrapper for ``R/src/training/utils.py:99-102``."""
    return manager.save(tag, engines, trainer_state, **kw)
