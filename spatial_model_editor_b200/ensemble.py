"""Parameter-scan ensembles (BASELINE config 5): independent simulations that
differ only in the values of runtime model parameters, sharded one simulation
per GPU slot — "replicas only": no data-path collective.

The reference has no scan feature: a scan is a Python loop that sets a
parameter and calls ``simulate()`` again, re-JIT-compiling the reactions every
time because parameter values are inlined into the expression text
(sme_model.cpp:541-617, pde.cpp:65-77; docs/sme/notebooks/simulating.ipynb).
Here the scanned parameters stay symbolic (``sme_param_k`` kernel arguments), so
every member reuses one NVRTC-compiled module, and the members of one GPU run
concurrently on their own CUDA streams, driven by a small pool of host threads
(the ctypes calls release the GIL).
"""
from __future__ import annotations

import copy
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

import ctypes as C

from .pixelsim import PixelSim, _STOP_CB, load_library
from .problem import Problem


def scan_values(centre: float, n_members: int, factor: float = 10.0,
                seed: int = 5000) -> np.ndarray:
    """log-uniform samples in [centre / factor, centre * factor] (SURVEY.md
    section 8(d), config C5: seed 5000)"""
    rng = np.random.default_rng(seed)
    return centre * factor ** rng.uniform(-1.0, 1.0, size=n_members)


def members_of_rank(n_members: int, rank: int, world: int) -> List[int]:
    """member m runs on rank m mod world"""
    return list(range(rank, n_members, world))


@dataclass
class MemberResult:
    member: int
    params: np.ndarray
    steps: int
    discarded: int
    error: str
    concentrations: List[np.ndarray]
    device_ms: float


class Ensemble:
    """``problem.params`` holds the base values of the runtime parameters;
    ``values[m]`` are member m's."""

    def __init__(self, problem: Problem, values: Sequence[Sequence[float]],
                 rank: int = 0, world: int = 1, device: int = 0, concurrency: int = 8,
                 sim_factory=None):
        self.problem = problem
        self.values = np.atleast_2d(np.asarray(values, dtype=np.float64))
        if self.values.shape[0] == 1 and len(problem.params) == 1 and \
                np.asarray(values).ndim == 1:
            self.values = self.values.T
        assert self.values.shape[1] == len(problem.params)
        self.rank, self.world, self.device = rank, world, device
        self.concurrency = max(1, concurrency)
        self.members = members_of_rank(self.values.shape[0], rank, world)
        self._factory = sim_factory or PixelSim

    def _make(self, m: int):
        p = copy.copy(self.problem)
        p.params = list(self.values[m])
        p.device = self.device
        return self._factory(p)

    def _run_member(self, m: int, time: float, sim=None) -> MemberResult:
        sim = sim or self._make(m)
        err = sim.errorMessage()
        steps = sim.run(time) if not err else 0
        res = MemberResult(
            member=m, params=self.values[m].copy(), steps=steps,
            discarded=sim.discardedSteps, error=err or sim.errorMessage(),
            concentrations=[sim.getConcentrations(ci).copy()
                            for ci in range(len(self.problem.compartments))],
            device_ms=sim.lastRunMs() if hasattr(sim, "lastRunMs") else 0.0)
        sim.close()
        return res

    def run(self, time: float) -> List[MemberResult]:
        """simulate this rank's members for ``time``; results in member order"""
        if not self.members:
            return []
        # the first member compiles the module; the others hit the cache
        first = self._run_member(self.members[0], time)
        rest = self.members[1:]
        if not rest:
            return [first]
        with ThreadPoolExecutor(max_workers=min(self.concurrency, len(rest))) as pool:
            others = list(pool.map(lambda m: self._run_member(m, time), rest))
        return [first] + others


class BatchedEnsemble:
    """All members of this rank in ONE simulator object behind the C-ABI
    (``smeb200_ensemble_*``): the state is [member][voxel][species], a Runge-Kutta
    stage of all members of a compartment is one kernel launch (members =
    gridDim.y), every member keeps its own step-size controller.  Same constructor
    and ``run`` result as ``Ensemble``; replaces its thread pool of launch-bound
    single simulations (40 us per step and member)."""

    def __init__(self, problem: Problem, values: Sequence[Sequence[float]],
                 rank: int = 0, world: int = 1, device: int = 0):
        self.problem = problem
        self.values = np.atleast_2d(np.asarray(values, dtype=np.float64))
        if self.values.shape[0] == 1 and len(problem.params) == 1 and \
                np.asarray(values).ndim == 1:
            self.values = self.values.T
        assert self.values.shape[1] == len(problem.params)
        self.members = members_of_rank(self.values.shape[0], rank, world)
        self._open(device)

    @classmethod
    def of_members(cls, problem: Problem, values: np.ndarray, members: List[int],
                   device: int = 0) -> "BatchedEnsemble":
        """an ensemble of an explicit list of member indices into ``values``"""
        self = cls.__new__(cls)
        self.problem, self.values, self.members = problem, values, list(members)
        self._open(device)
        return self

    def _open(self, device: int) -> None:
        problem = self.problem
        self._lib = load_library()
        self._h = C.c_void_p()
        p = copy.copy(problem)
        p.device = device
        cproblem, keep = p.to_c()
        mine = np.ascontiguousarray(self.values[self.members])
        self._lib.smeb200_ensemble_create(
            C.byref(cproblem), len(self.members),
            mine.ctypes.data_as(C.POINTER(C.c_double)), C.byref(self._h))
        del keep
        if not self._h:
            raise RuntimeError("ensemble construction failed")

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.smeb200_ensemble_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def errorMessage(self, member: int = -1) -> str:
        return self._lib.smeb200_ensemble_error(self._h, member).decode()

    def setNextTimestep(self, dt: float) -> None:
        for k in range(len(self.members)):
            self._lib.smeb200_ensemble_set_next_timestep(self._h, k, dt)

    def launchCount(self) -> int:
        return int(self._lib.smeb200_ensemble_launch_count(self._h))

    def batchedSteps(self) -> int:
        return int(self._lib.smeb200_ensemble_batched_steps(self._h))

    def lastRunMs(self) -> float:
        return float(self._lib.smeb200_ensemble_last_run_ms(self._h))

    def getConcentrations(self, k: int, compartmentIndex: int) -> np.ndarray:
        p = C.POINTER(C.c_double)()
        n = C.c_size_t()
        rc = self._lib.smeb200_ensemble_get_concentrations(
            self._h, k, compartmentIndex, C.byref(p), C.byref(n))
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        S = max(self.problem.compartments[compartmentIndex].n_species, 1)
        if n.value == 0:
            return np.zeros((0, S))
        return np.ctypeslib.as_array(p, shape=(n.value,)).reshape(-1, S)

    def run(self, time: float, timeout_ms: float = -1.0) -> List[MemberResult]:
        if self.errorMessage():
            raise RuntimeError(self.errorMessage())
        self._lib.smeb200_ensemble_run(self._h, time, timeout_ms, _STOP_CB(), None)
        ncomp = len(self.problem.compartments)
        out = []
        for k, m in enumerate(self.members):
            out.append(MemberResult(
                member=m, params=self.values[m].copy(),
                steps=int(self._lib.smeb200_ensemble_get_steps(self._h, k)),
                discarded=int(self._lib.smeb200_ensemble_get_discarded_steps(self._h, k)),
                error=self.errorMessage(k) or self.errorMessage(),
                concentrations=[self.getConcentrations(k, ci).copy()
                                for ci in range(ncomp)],
                device_ms=self.lastRunMs()))
        return out


class GroupedEnsemble:
    """This rank's members as ``groups`` member-batched ensembles driven by one
    host thread each: while one group's step-size controllers look at the error
    norms of its step, the other group's step is running — the GPU never waits for
    the host round trip between two batched steps."""

    def __init__(self, problem: Problem, values: Sequence[Sequence[float]],
                 rank: int = 0, world: int = 1, device: int = 0, groups: int = 2):
        probe = BatchedEnsemble.__new__(BatchedEnsemble)
        v = np.atleast_2d(np.asarray(values, dtype=np.float64))
        if v.shape[0] == 1 and len(problem.params) == 1 and np.asarray(values).ndim == 1:
            v = v.T
        self.values = v
        mine = members_of_rank(v.shape[0], rank, world)
        groups = max(1, min(groups, len(mine)))
        self.parts = [BatchedEnsemble.of_members(problem, v, mine[g::groups], device)
                      for g in range(groups)]
        del probe

    def errorMessage(self) -> str:
        return "; ".join(e for e in (p.errorMessage() for p in self.parts) if e)

    def launchCount(self) -> int:
        return sum(p.launchCount() for p in self.parts)

    def batchedSteps(self) -> int:
        return sum(p.batchedSteps() for p in self.parts)

    def run(self, time: float) -> List[MemberResult]:
        if len(self.parts) == 1:
            res = self.parts[0].run(time)
        else:
            with ThreadPoolExecutor(max_workers=len(self.parts)) as pool:
                res = [r for part in pool.map(lambda p: p.run(time), self.parts)
                       for r in part]
        return sorted(res, key=lambda r: r.member)

    def close(self) -> None:
        for p in self.parts:
            p.close()
