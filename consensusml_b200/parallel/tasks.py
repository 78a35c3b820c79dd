"""Static task scheduler: spread the independent pieces of a pipeline run over the ranks.

The reference runs its ensemble strictly one model after the other in one R session (cml_targetaml_seanalysis.Rmd:638-1298:
SVM x4 at :648-667, 15 x LOOCV lasso at :728-773, RF 2k/5k/10k at :1037-1050, xgboost x5 at :1164-1186).  None of those
stages depends on another - only the consensus table at the end needs all of them - so on N GPUs they are independent
tasks: the sequential lasso exclusion chain stays on one rank, the 17,000 forest trees are cut into fixed 1000-tree
chunks, every SVM / boosting / MLP configuration is a task of its own.

Design rules
  * The task list and its chunk boundaries never depend on the world size, and every task derives its random streams
    from its own identity (tree index, replicate index), so the merged result is the same for 1, 2, 4 or 8 ranks.
  * Assignment is longest-processing-time-first on declared costs, computed identically on every rank (no
    communication, no work stealing: the run is a second long, a dynamic queue would cost more than it saves).
  * Results are small (per-gene vectors, n x n proximity sums); one ``all_gather_object`` at the end gives every rank
    every result.  Device time per task is recorded so the limiting task can be named.
  * Stage overlap inside a rank (``streams > 1``): the rank's tasks run on a small pool of host threads, each with its own
    CUDA stream, heaviest first.  The sequential lasso chain is latency bound (one CTA per fit, at most 94 of 148 SMs, long
    dependent launches) - the forest chunks, SVMs and boosting runs fill the rest of the machine meanwhile.  Tasks only read
    the shared cohort tensors and return host results, so they need no synchronisation among themselves.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional

import torch
import torch.distributed as dist


@dataclass
class Task:
    name: str
    cost: float
    fn: Callable[[], Any]
    group: str = ""


@dataclass
class TaskReport:
    assignment: Dict[str, int] = field(default_factory=dict)     # task -> rank
    seconds: Dict[str, float] = field(default_factory=dict)      # task -> wall seconds on its rank
    rank_seconds: List[float] = field(default_factory=list)      # per rank: sum of its tasks

    def critical(self) -> Dict[str, Any]:
        if not self.rank_seconds:
            return {}
        r = max(range(len(self.rank_seconds)), key=lambda i: self.rank_seconds[i])
        mine = {k: v for k, v in self.seconds.items() if self.assignment.get(k) == r}
        top = max(mine, key=mine.get) if mine else None
        return {"rank": r, "seconds": self.rank_seconds[r], "largest_task": top, "largest_task_seconds": mine.get(top)}


def assign(tasks: List[Task], world: int) -> Dict[str, int]:
    """Longest-processing-time-first: heaviest task to the least loaded rank (ties: lowest rank, input order)."""
    load = [0.0] * world
    out: Dict[str, int] = {}
    order = sorted(range(len(tasks)), key=lambda i: (-tasks[i].cost, i))
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[tasks[i].name] = r
        load[r] += tasks[i].cost
    return out


class TaskError:
    """Result of a task that raised (``fail_fast=False``): the ensemble goes on without that stage and reports it."""

    def __init__(self, name: str, exc: BaseException):
        self.name, self.error = name, f"{type(exc).__name__}: {exc}"

    def __repr__(self):
        return f"TaskError({self.name!r}, {self.error!r})"


def _run_one(t: Task, fail_fast: bool, stream) -> tuple:
    """(result, seconds) of one task, run on ``stream`` when given (device work is complete on return)."""
    t0 = time.perf_counter()
    try:
        if stream is not None:
            with torch.cuda.stream(stream):
                out = t.fn()
            stream.synchronize()
        else:
            out = t.fn()
            if torch.cuda.is_available():
                torch.cuda.synchronize()
    except Exception as exc:                           # noqa: BLE001 - reported, not swallowed
        if fail_fast:
            raise
        out = TaskError(t.name, exc)
    return out, time.perf_counter() - t0


def _run_overlapped(mine: List[Task], streams: int, fail_fast: bool, local: Dict[str, Any], secs: Dict[str, float]) -> None:
    import threading
    from concurrent.futures import ThreadPoolExecutor
    use_cuda = torch.cuda.is_available()
    tls = threading.local()
    main = torch.cuda.current_stream() if use_cuda else None
    made: List[Any] = []
    lock = threading.Lock()
    dev = torch.cuda.current_device() if use_cuda else None

    def work(t: Task):
        if use_cuda and not hasattr(tls, "stream"):
            torch.cuda.set_device(dev)                 # a new host thread starts on device 0
            tls.stream = torch.cuda.Stream()
            tls.stream.wait_stream(main)               # everything queued before the stage started is visible
            with lock:
                made.append(tls.stream)
        return _run_one(t, fail_fast, getattr(tls, "stream", None))

    order = sorted(range(len(mine)), key=lambda i: (-mine[i].cost, i))
    with ThreadPoolExecutor(max_workers=streams, thread_name_prefix="cml-task") as pool:
        futs = {mine[i].name: pool.submit(work, mine[i]) for i in order}
        err = None
        for name, f in futs.items():
            try:
                local[name], secs[name] = f.result()
            except Exception as exc:                   # noqa: BLE001 - first failure is re-raised once every thread has finished
                err = err or exc
    for s in made:
        main.wait_stream(s)
    if err is not None:
        raise err


def run_tasks(tasks: List[Task], *, enabled: bool = True, fail_fast: bool = True, streams: int = 1) -> tuple[Dict[str, Any], TaskReport]:
    """Run every task on the rank it is assigned to and return ``{task name: result}`` complete on every rank.
    ``streams > 1``: this rank's tasks overlap on that many CUDA streams (host threads); ``seconds`` are then wall times of
    overlapping intervals and ``rank_seconds`` the wall time of the whole stage.
    Results must be picklable and live on the host (tasks copy what they return off the device).
    ``fail_fast=False`` turns an exception inside a task into a ``TaskError`` result instead of aborting the run (every rank
    still reaches the gather, so a failed stage cannot leave the other ranks waiting)."""
    names = [t.name for t in tasks]
    if len(set(names)) != len(names):
        raise ValueError("task names must be unique")
    world = dist.get_world_size() if (enabled and dist.is_initialized()) else 1
    rank = dist.get_rank() if world > 1 else 0
    amap = assign(tasks, world)
    local: Dict[str, Any] = {}
    secs: Dict[str, float] = {}
    mine = [t for t in tasks if amap[t.name] == rank]
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    stage0 = time.perf_counter()
    from .. import _native
    if streams > 1 and len(mine) > 1 and not _native.emulating():      # the host execution model is single-threaded by design
        _run_overlapped(mine, int(streams), fail_fast, local, secs)
    else:
        for t in mine:
            local[t.name], secs[t.name] = _run_one(t, fail_fast, None)
    stage_s = time.perf_counter() - stage0
    if world > 1:
        gathered: List[Optional[tuple]] = [None] * world
        dist.all_gather_object(gathered, (local, secs, stage_s))
        results: Dict[str, Any] = {}
        all_secs: Dict[str, float] = {}
        stage = []
        for part, s, st in gathered:
            results.update(part)
            all_secs.update(s)
            stage.append(st)
    else:
        results, all_secs, stage = local, secs, [stage_s]
    rank_s = stage if streams > 1 else [sum(v for k, v in all_secs.items() if amap[k] == r) for r in range(world)]
    rep = TaskReport(amap, all_secs, rank_s)
    return {n: results[n] for n in names}, rep
