"""One configuration tree for the whole pipeline.

The reference configures itself with literals scattered through its notebooks: file-name globals
(cml_targetaml_seanalysis.Rmd:370-407; seobjects/make_seobj_targetaml.R:21-36), inline hyper-parameters
(scripts/model_comp.py:7-9; cml_targetaml_seanalysis.Rmd:92-101,1164-1186; VikasP/AML.R:241-245), thresholds
(JSmith_code/Limma_Voom_DE_Function.R:27,42) and seeds 2019 / 50 / 20 / 8 / 2128 / 1000.  The defaults below
reproduce those literals; every field can be overridden from YAML / JSON or ``section.key=value`` CLI arguments,
and the hash of the resolved tree is written into every artefact."""
from __future__ import annotations

import dataclasses
import hashlib
import json
from dataclasses import asdict, dataclass, field
from typing import Any, Dict, List, Optional, Sequence


@dataclass
class IOConfig:
    counts_csv: Optional[str] = None
    clinical_csv: Optional[str] = None
    cohort: str = "target_aml"            # bundled fixture when no files are given
    out_dir: str = "cml_out"


@dataclass
class NormalizeConfig:
    min_cpm: float = 1.0
    min_samples: float = 5                 # DEL:221 (NRM:55 uses max(2, 0.05 n))
    prior_count: float = 1.0


@dataclass
class DEConfig:
    min_frac: float = 0.05                 # VDE:27
    p_value: float = 0.05                  # VDE:42
    lfc: float = 1.0
    span: float = 0.5


@dataclass
class SplitConfig:
    mode: str = "fixed"                    # fixed | random | stratified
    train_frac: float = 2.0 / 3.0
    seed: int = 2019


@dataclass
class LassoConfig:
    enabled: bool = True
    alpha: float = 1.0
    nfolds: Optional[int] = None           # None = leave-one-out (SEA:100)
    exclusion_reps: int = 15              # SEA:728-773 ran 15 replicates
    seed: int = 2019
    stability_boot: int = 0                # > 0: stability selection with this many bootstrap replicates per alpha (BASELINE config 5)
    stability_alphas: List[float] = field(default_factory=lambda: [1.0, 0.5])
    stability_chunk: int = 250             # bootstraps per scheduler task (fixed: results do not depend on the world size)
    stability_threshold: float = 0.6       # a gene is "stable" when its maximal selection frequency reaches this


@dataclass
class SVMConfig:
    enabled: bool = True
    kernels: List[str] = field(default_factory=lambda: ["linear", "radial"])
    weightfilt: List[Optional[float]] = field(default_factory=lambda: [None, 0.5])
    seed: int = 50


@dataclass
class RFConfig:
    enabled: bool = True
    n_trees: List[int] = field(default_factory=lambda: [2000, 5000, 10000])
    proximity: bool = True
    seed: int = 20
    chunk: int = 1000                      # trees per scheduler task; fixed so that results do not depend on the world size


@dataclass
class GBMConfig:
    enabled: bool = True
    reps: List[Dict[str, float]] = field(default_factory=lambda: [
        {"max_depth": 2, "nrounds": 2}, {"max_depth": 50, "nrounds": 2}, {"max_depth": 50, "nrounds": 50},
        {"max_depth": 100, "nrounds": 50}, {"max_depth": 100, "nrounds": 100}])      # SEA:1164-1186
    eta: float = 1.0


@dataclass
class MLPConfig:
    enabled: bool = False
    hidden: List[int] = field(default_factory=lambda: [16, 32, 64])
    lr: List[float] = field(default_factory=lambda: [1e-3, 3e-3])
    weight_decay: List[float] = field(default_factory=lambda: [0.0, 1e-3])
    epochs: int = 200
    val_frac: float = 0.25                 # inner split of the training samples used to pick the grid point
    seed: int = 0


@dataclass
class ConsensusConfig:
    method: str = "rra"                    # rra | mean_rank | borda
    top: int = 50


@dataclass
class DistConfig:
    enabled: bool = True                   # under torchrun: spread the independent stages over the ranks (parallel/tasks.py)
    backend: str = "auto"                  # nccl on GPUs, gloo on CPU (parallel.dist.init)
    fused: bool = True                     # large sample-sharded cohorts: in-kernel collectives (models/lasso_large, gbm_large)
    task_streams: int = 1                  # > 1: the stages a rank owns run concurrently, each on its own CUDA stream (opt-in: not yet timed on hardware)


@dataclass
class Config:
    io: IOConfig = field(default_factory=IOConfig)
    normalize: NormalizeConfig = field(default_factory=NormalizeConfig)
    de: DEConfig = field(default_factory=DEConfig)
    split: SplitConfig = field(default_factory=SplitConfig)
    lasso: LassoConfig = field(default_factory=LassoConfig)
    svm: SVMConfig = field(default_factory=SVMConfig)
    rf: RFConfig = field(default_factory=RFConfig)
    gbm: GBMConfig = field(default_factory=GBMConfig)
    mlp: MLPConfig = field(default_factory=MLPConfig)
    consensus: ConsensusConfig = field(default_factory=ConsensusConfig)
    dist: DistConfig = field(default_factory=DistConfig)
    device: str = "auto"
    resume: bool = True
    fail_fast: bool = True                 # False: a stage that raises is reported in summary["failed_tasks"], the others go on
    checkpoint: bool = True                # write per-task artefacts (and reuse them when `resume`)
    write_outputs: bool = True             # standouttable / gene CSV / report dumps (benchmarks of the compute switch this off)

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    def hash(self, sections: Optional[Sequence[str]] = None) -> str:
        d = self.to_dict()
        if sections is not None:
            d = {k: d[k] for k in sections}
        return hashlib.sha256(json.dumps(d, sort_keys=True, default=str).encode()).hexdigest()[:16]


def _coerce(old, text: str):
    if isinstance(old, bool):
        return text.lower() in ("1", "true", "yes", "on")
    if isinstance(old, int) and not isinstance(old, bool):
        return int(text)
    if isinstance(old, float):
        return float(text)
    if isinstance(old, (list, dict)) or old is None:
        try:
            return json.loads(text)
        except json.JSONDecodeError:
            return text
    return text


def apply_overrides(cfg: Config, overrides: Sequence[str]) -> Config:
    """``section.key=value`` (or ``key=value`` for top-level fields)."""
    for ov in overrides:
        if "=" not in ov:
            raise ValueError(f"override '{ov}' is not of the form section.key=value")
        key, text = ov.split("=", 1)
        obj = cfg
        parts = key.split(".")
        for p in parts[:-1]:
            if not hasattr(obj, p):
                raise KeyError(f"unknown config section '{p}' in '{ov}'")
            obj = getattr(obj, p)
        if not hasattr(obj, parts[-1]):
            raise KeyError(f"unknown config key '{key}'")
        setattr(obj, parts[-1], _coerce(getattr(obj, parts[-1]), text))
    return cfg


def load(path: Optional[str] = None, overrides: Sequence[str] = ()) -> Config:
    cfg = Config()
    if path:
        with open(path) as fh:
            if path.endswith((".yaml", ".yml")):
                import yaml
                data = yaml.safe_load(fh) or {}
            else:
                data = json.load(fh)
        for sec, vals in data.items():
            if dataclasses.is_dataclass(getattr(cfg, sec, None)):
                for k, v in vals.items():
                    if not hasattr(getattr(cfg, sec), k):
                        raise KeyError(f"unknown config key '{sec}.{k}'")
                    setattr(getattr(cfg, sec), k, v)
            elif hasattr(cfg, sec):
                setattr(cfg, sec, vals)
            else:
                raise KeyError(f"unknown config section '{sec}'")
    return apply_overrides(cfg, overrides)
