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
import time
from pathlib import Path
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from dla_b200.models.io import SafetensorsReader, ShardedSafetensorsWriter, export_unit, has_pretrained_weights, import_unit
from dla_b200.parallel.engine import Engine
from dla_b200.parallel.pg import DistContext


def _snapshot_model(engine: Engine, ctx: DistContext, reward_bin: bool = False):
    """Collective part of a model export: every rank takes part in the unit gathers, rank 0 ends up with the
    HF-named tensors on the HOST (and the ``backbone.*`` / ``scorer.*`` state of the reward phase)."""
    tensors: Dict[str, torch.Tensor] = {}
    bin_state: Dict[str, torch.Tensor] = {}
    for u in engine.units:
        w = engine.gathered_unit(u.name)
        if not ctx.is_main:
            continue
        for k, t in export_unit(engine.cfg, u.name, w).items():
            t = t.detach().to("cpu").contiguous()
            tensors[k] = t
            if reward_bin:
                if k.startswith("scorer."):
                    bin_state[k] = t
                else:
                    kk = k.split(".", 1)[1] if k.split(".", 1)[0] in ("model", "transformer") else k
                    bin_state["backbone." + kk] = t
    return tensors, bin_state


def _write_model(engine_cfg: Any, out_dir: Path, tensors: Dict[str, torch.Tensor], bin_state: Dict[str, torch.Tensor],
                 tokenizer: Any, reward_bin: bool) -> None:
    """File part (rank 0 only, no collectives): may run on the writer thread."""
    writer = ShardedSafetensorsWriter(out_dir)
    writer.add(tensors)
    writer.close()
    engine_cfg.save(out_dir)
    if tokenizer is not None and hasattr(tokenizer, "save_pretrained"):
        tokenizer.save_pretrained(str(out_dir))
    if reward_bin:
        torch.save(bin_state, out_dir / "pytorch_model.bin")


def save_model_dir(engine: Engine, out_dir: Path, ctx: DistContext, tokenizer: Any = None,
                   reward_bin: bool = False) -> None:
    """Consolidated HF-style model directory.  Collective: every rank takes part in the unit gathers,
    rank 0 writes."""
    tensors, bin_state = _snapshot_model(engine, ctx, reward_bin)
    if ctx.is_main:
        _write_model(engine.cfg, Path(out_dir), tensors, bin_state, tokenizer, reward_bin)


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
    engine.publish_shards()     # streamed fused engines: peers read these shards directly, make them visible first
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
    """``async_write=True`` splits a save into a snapshot (collective gathers + device -> host copies, done before
    ``save`` returns, so training may overwrite the weights right away) and a write phase on a background thread
    (serialisation, fsync-free file IO, the atomic rename, ``latest`` and pruning).  The thread uses no collective:
    ranks signal completion with marker files and rank 0 finalises once every rank's marker exists.  ``wait()`` (also
    called by the next ``save`` and by ``load``) joins it and re-raises its error."""

    def __init__(self, output_dir: str | Path, ctx: DistContext, keep_last_n: Optional[int] = None,
                 async_write: bool = False, finalize_timeout_s: float = 1800.0) -> None:
        self.root, self.ctx, self.keep = Path(output_dir), ctx, keep_last_n
        self.async_write = async_write
        self.finalize_timeout_s = finalize_timeout_s
        self._thread: Optional[threading.Thread] = None
        self._error: Optional[BaseException] = None
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
        # ---- snapshot: everything that needs the device or a collective ---------------------------
        primary = primary or next((n for n, e in engines.items() if e.trainable), next(iter(engines)))
        model_tensors, bin_state = _snapshot_model(engines[primary], self.ctx, reward_bin)
        shard_states: Dict[str, Dict[str, torch.Tensor]] = {}
        rng = None
        if save_optimizer:
            for name, eng in engines.items():
                if eng.trainable:
                    shard_states[name] = {k: v.detach().cpu().contiguous().clone() if v.device.type == "cpu"
                                          else v.detach().cpu().contiguous() for k, v in eng.shard_state().items()}
            rng = _rng_state()
        meta = {"world_size": self.ctx.world_size,
                "engines": {n: {"sharding": e.sharding, "step_count": e.step_count, "trainable": e.trainable,
                                "shard_group": e.W} for n, e in engines.items()}}
        trainer_state = json.loads(json.dumps(trainer_state, default=str))      # frozen copy for the writer
        cfg = engines[primary].cfg
        rank, world, is_main = self.ctx.rank, self.ctx.world_size, self.ctx.is_main

        # ---- write: files only, no collectives ------------------------------------------------------
        def write() -> None:
            from safetensors.torch import save_file

            for name, state in shard_states.items():
                save_file(state, str(tmp / "shards" / f"rank_{rank}.{name}.safetensors"))
            if rng is not None:
                torch.save(rng, tmp / "shards" / f"rng_{rank}.pt")
            if is_main:
                _write_model(cfg, tmp, model_tensors, bin_state, tokenizer, reward_bin)
            (tmp / "shards" / f".done_{rank}").touch()
            if not is_main:
                return
            deadline = time.monotonic() + self.finalize_timeout_s
            while not all((tmp / "shards" / f".done_{r}").exists() for r in range(world)):
                if time.monotonic() > deadline:
                    raise TimeoutError(f"checkpoint {tag}: not every rank finished writing within {self.finalize_timeout_s}s")
                time.sleep(0.02)
            for r in range(world):
                (tmp / "shards" / f".done_{r}").unlink()
            with (tmp / "shards" / "meta.json").open("w") as fh:
                json.dump(meta, fh, indent=2)
            with (tmp / "trainer_state.json").open("w") as fh:
                json.dump(trainer_state, fh, indent=2)
            if out.exists():
                shutil.rmtree(out)
            os.replace(tmp, out)
            self._point_latest(tag)
            self._prune()

        if self.async_write:
            def guarded() -> None:
                try:
                    write()
                except BaseException as e:  # surfaced by wait()
                    self._error = e

            self._thread = threading.Thread(target=guarded, name=f"ckpt-writer-{tag}", daemon=False)
            self._thread.start()
        else:
            write()
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
        """Block until a background write (if any) has finished; raises what it raised."""
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        if self._error is not None:
            e, self._error = self._error, None
            raise RuntimeError(f"background checkpoint write failed: {e!r}") from e

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

        self.wait()
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
