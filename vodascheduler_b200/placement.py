"""Host-side mirror of the reference's placement interface.

    m = munkres.NewMatrix(n); m.A = [...]; munkres.ComputeMunkresMax(m)   -> [RowCol(Row, Col)]
    pm = PlacementManager(nodes); pm.Place(jobRequests)                    -> pods to delete

`Place` follows pkg/placement/placement_manager.go:306-332.  The arithmetic core — bestFit
(:415-487), the overlap score matrix (:525-544), ComputeMunkresMax (:505-507) and the rename
(:509-511) — runs on the GPU through voda_place; the bookkeeping around it (releaseSlots
:337-411, updateJobStates :548-566, updatePodNodeName :571-617) is plain dictionary work and
stays on the host, restated line by line.  Everything the Go code takes from map iteration
order uses Python's insertion order here, which makes the results deterministic (the
reference's are not).  The k8s half of the manager (informers, tolerations, pod
deletion) is out of scope.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, NamedTuple

import numpy as np

from .batch import PlaceBatch


class RowCol(NamedTuple):
    Row: int
    Col: int


class Matrix:
    """munkres.Matrix: n and the row-major A []int64 (index n*row + col)."""

    def __init__(self, n: int):
        self.n = n
        self.A = [0] * (n * n)


class munkres:  # namespace mirroring the Go package name used at placement_manager.go:505-507
    _ctx = None

    @staticmethod
    def NewMatrix(n: int) -> Matrix:
        return Matrix(n)

    @staticmethod
    def ComputeMunkresMax(m: Matrix, ctx=None) -> List[RowCol]:
        if ctx is None:
            if munkres._ctx is None:
                from ._lib import Context
                munkres._ctx = Context(0)
            ctx = munkres._ctx
        if m.n == 0:
            return []
        cols, st = ctx.munkres_max(np.asarray(m.A, np.int64).reshape(1, m.n, m.n))
        if st[0] != 0:
            from ._lib import status_str
            raise RuntimeError(status_str(st[0]))
        return [RowCol(r, int(c)) for r, c in enumerate(cols[0])]


@dataclass
class nodeNumSlots:  # pkg/placement/types.go:16-19
    node: str
    numSlots: int


@dataclass
class jobState:  # types.go:22-37
    name: str
    numWorkers: int = 0
    nodeNumSlotsList: List[nodeNumSlots] = field(default_factory=list)


@dataclass
class nodeState:  # types.go:39-50
    name: str
    totalSlots: int
    freeSlots: int
    jobNumWorkers: Dict[str, int] = field(default_factory=dict)


def NewNodeState(name: str, slots: int) -> nodeState:
    return nodeState(name, slots, slots, {})


def getWorkerPodName(job: str, idx: int) -> str:  # pkg/placement/utils.go
    return f"{job}-worker-{idx}"


def getLauncherPodName(job: str) -> str:
    return f"{job}-launcher"


class PlacementError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


class PlacementManager:
    """Arithmetic state of the reference's PlacementManager (nodeStates, jobStates,
    podNodeName); no Kubernetes objects."""

    def __init__(self, nodes: Dict[str, int], ctx=None):
        self.nodeStates: Dict[str, nodeState] = {n: NewNodeState(n, s) for n, s in nodes.items()}
        self.jobStates: Dict[str, jobState] = {}
        self.podNodeName: Dict[str, str] = {}
        self._ctx = ctx
        self.lastScore = 0
        self.lastCrossNode = 0
        self.deletedWorkers = 0
        self.deletedLaunchers = 0

    def _context(self):
        if self._ctx is None:
            from ._lib import Context
            self._ctx = Context(0)
        return self._ctx

    # placement_manager.go:337-411
    def releaseSlots(self, jobRequests: Dict[str, int]) -> None:
        for job in self.jobStates.values():
            if job.name not in jobRequests:
                for nSlots in job.nodeNumSlotsList:
                    node = self.nodeStates.get(nSlots.node)
                    if node is not None:
                        node.freeSlots += nSlots.numSlots
                        node.jobNumWorkers.pop(job.name, None)
                job.nodeNumSlotsList = []
            elif jobRequests[job.name] < job.numWorkers:
                toRelease = job.numWorkers - jobRequests[job.name]
                while toRelease > 0:
                    last = job.nodeNumSlotsList[-1]
                    node = self.nodeStates.get(last.node)
                    if last.numSlots >= toRelease:
                        last.numSlots -= toRelease
                        node.freeSlots += toRelease
                        node.jobNumWorkers[job.name] = node.jobNumWorkers.get(job.name, 0) - toRelease
                        toRelease = 0
                    else:
                        toRelease -= last.numSlots
                        last.numSlots = 0
                        if node is not None:
                            node.freeSlots += node.jobNumWorkers.get(job.name, 0)
                            node.jobNumWorkers[job.name] = 0
                    if last.numSlots == 0:
                        job.nodeNumSlotsList.pop()
                    if node is not None and node.jobNumWorkers.get(job.name, 0) == 0:
                        node.jobNumWorkers.pop(job.name, None)

    def pack(self, jobRequests: Dict[str, int]) -> PlaceBatch:
        """One placement round as a PlaceBatch (explicit orders = insertion orders)."""
        names = list(jobRequests)
        idx = {n: i for i, n in enumerate(names)}
        nodes = list(self.nodeStates.values())
        prev_off, pj, pw = [0], [], []
        for nd in nodes:
            for job, w in nd.jobNumWorkers.items():
                if job in idx:
                    pj.append(idx[job])
                    pw.append(w)
            prev_off.append(len(pj))
        return PlaceBatch([0, len(nodes)], [nd.totalSlots for nd in nodes], [0, len(names)],
                          [jobRequests[n] for n in names], prev_off, np.array(pj, np.int64),
                          np.array(pw, np.int64))

    # placement_manager.go:306-332
    def Place(self, jobRequests: Dict[str, int]) -> List[str]:
        batch = self.begin_place(jobRequests)
        out = self._context().place(batch)
        return self.finish_place(jobRequests, out, 0)

    def begin_place(self, jobRequests: Dict[str, int]) -> PlaceBatch:
        """releaseSlots + the round as a PlaceBatch (first half of Place)."""
        self.releaseSlots(jobRequests)
        return self.pack(jobRequests)

    def finish_place(self, jobRequests: Dict[str, int], out, rnd: int, node_base: int = 0,
                     entry_base: int = 0) -> List[str]:
        """Second half of Place: apply round `rnd` of a (possibly batched) voda_place result.
        node_base / entry_base locate the round inside the batch's flat output arrays."""
        if out.status[rnd] != 0:
            from ._lib import status_str
            raise PlacementError(int(out.status[rnd]), status_str(out.status[rnd]))
        names = list(jobRequests)
        current = list(self.nodeStates.values())  # currentNodeList (:494-497)
        n = len(current)
        # logical nodes filled by bestFit, then renamed to the physical node they bind to
        logical = [NewNodeState("TBD", nd.totalSlots) for nd in current]
        for t in range(entry_base, entry_base + int(out.lay_cnt[rnd])):
            logical[int(out.lay_node[t])].jobNumWorkers[names[int(out.lay_job[t])]] = int(out.lay_workers[t])
        for i in range(n):
            logical[i].freeSlots = int(out.free_slots[node_base + i])
            logical[i].name = current[int(out.phys_of_logical[node_base + i])].name  # rename (:509-511)
        self.lastScore = int(out.score[rnd])
        self.lastCrossNode = int(out.cross[rnd])
        # Go keeps nodeStates in a map (random iteration order); here the physical nodes keep their
        # first order for ever, the defined order of include/voda_accel.h ("trace replay")
        by_name = {nd.name: nd for nd in logical}
        self.nodeStates = {cur.name: by_name[cur.name] for cur in current}
        self.updateJobStates()
        return self.updatePodNodeName()

    # placement_manager.go:548-566
    def updateJobStates(self) -> None:
        new: Dict[str, jobState] = {}
        for node in self.nodeStates.values():
            for jobName, workers in node.jobNumWorkers.items():
                js = new.setdefault(jobName, jobState(jobName))
                js.nodeNumSlotsList.append(nodeNumSlots(node.name, workers))
                js.numWorkers += workers
        self.jobStates = new

    # placement_manager.go:571-617
    def updatePodNodeName(self) -> List[str]:
        newPodNodeName: Dict[str, str] = {}
        deleting: List[str] = []
        self.deletedWorkers = self.deletedLaunchers = 0
        for job in self.jobStates.values():
            idx = deleted = 0
            for nSlots in job.nodeNumSlotsList:
                for _ in range(nSlots.numSlots):
                    pod = getWorkerPodName(job.name, idx)
                    old = self.podNodeName.get(pod)
                    if old is not None and nSlots.node != old:
                        deleting.append(pod)
                        deleted += 1
                        self.deletedWorkers += 1
                    newPodNodeName[pod] = nSlots.node
                    idx += 1
            if deleted == job.numWorkers:
                deleting.append(getLauncherPodName(job.name))
                self.deletedLaunchers += 1
        self.podNodeName = newPodNodeName
        return deleting
