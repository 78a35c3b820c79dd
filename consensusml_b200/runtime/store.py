"""Per-stage artefact store with resume.

The reference checkpoints with R ``save()`` / ``load()`` between notebook chunks so that expensive fits are not
re-run (cml_targetaml_seanalysis.Rmd:472,673,778,1069,1239; seobjects/make_seobj_targetaml.R:227-229).  Here every
pipeline stage writes ``<out>/<stage>/`` with its tensors (``.pt``), tables (``.csv``) and a ``manifest.json`` holding
the configuration hash of the sections it depends on, its seeds, timing and status; a re-run with ``resume`` skips a
stage whose manifest matches."""
from __future__ import annotations

import json
import os
import time
from typing import Any, Callable, Dict, Optional

import torch


class ArtifactStore:
    def __init__(self, root: str, resume: bool = True, rank: int = 0):
        self.root = root
        self.resume = resume
        self.rank = rank
        if rank == 0:
            os.makedirs(root, exist_ok=True)

    def dir(self, stage: str) -> str:
        d = os.path.join(self.root, stage)
        os.makedirs(d, exist_ok=True)
        return d

    def manifest_path(self, stage: str) -> str:
        return os.path.join(self.root, stage, "manifest.json")

    def is_done(self, stage: str, config_hash: str) -> bool:
        p = self.manifest_path(stage)
        if not (self.resume and os.path.exists(p)):
            return False
        try:
            m = json.load(open(p))
        except (OSError, json.JSONDecodeError):
            return False
        return m.get("status") == "ok" and m.get("config_hash") == config_hash

    def save(self, stage: str, name: str, obj: Any) -> str:
        path = os.path.join(self.dir(stage), name)
        tmp = path + ".tmp"
        torch.save(obj, tmp)
        os.replace(tmp, path)                 # atomic: a killed run never leaves a half-written checkpoint
        return path

    def load(self, stage: str, name: str) -> Any:
        return torch.load(os.path.join(self.root, stage, name), weights_only=False)

    def finish(self, stage: str, config_hash: str, **info) -> None:
        m = {"stage": stage, "status": "ok", "config_hash": config_hash, "finished": time.time(), **info}
        tmp = self.manifest_path(stage) + ".tmp"
        os.makedirs(os.path.dirname(tmp), exist_ok=True)
        with open(tmp, "w") as fh:
            json.dump(m, fh, indent=1, default=str)
        os.replace(tmp, self.manifest_path(stage))

    def run(self, stage: str, config_hash: str, fn: Callable[[], Dict[str, Any]], log=None) -> Dict[str, Any]:
        """Run ``fn`` unless a matching checkpoint exists; ``fn`` returns a dict of picklable results."""
        if self.is_done(stage, config_hash):
            if log:
                log.event("stage_skipped", stage=stage, reason="checkpoint matches configuration and data", key=config_hash)
            return self.load(stage, "result.pt")
        t0 = time.perf_counter()
        out = fn()
        self.save(stage, "result.pt", out)
        self.finish(stage, config_hash, seconds=time.perf_counter() - t0)
        return out
