"""CPU restatement of the placement manager's state bookkeeping — TEST INFRASTRUCTURE, not product.

Follows pkg/placement/placement_manager.go of the reference, statement by statement:
    releaseSlots        :337-411
    bindNodes (rename)  :492-522   (the score matrix and the assignment come from oracle.place)
    updateJobStates     :548-566
    updatePodNodeName   :571-617
written from the Go text independently of vodascheduler_b200/placement.py and of csrc/replay.cu,
which it checks.  State is kept the way the reference keeps it (dictionaries keyed by names).
Where the reference ranges over a Go map (random order) the order documented in
include/voda_accel.h is used: nodes by physical index, a node's jobs in insertion order.

Parity: the reference has no test for these functions and cannot be run here (no Go toolchain), so
this restatement is pinned by the hand-traced cases of tests/test_bookkeeping_cpu.py only:
"parity unpinned" against running Go.
"""
from __future__ import annotations


class NodeState:  # pkg/placement/types.go:27-46
    def __init__(self, name, total):
        self.name = name
        self.totalSlots = total
        self.freeSlots = total
        self.jobNumWorkers = {}  # job -> workers (dict keeps insertion order)


class JobState:  # pkg/placement/types.go:48-64
    def __init__(self, name):
        self.name = name
        self.numWorkers = 0
        self.nodeNumSlotsList = []  # [node, numSlots] pairs


class Manager:
    """nodeStates / jobStates / podNodeName of one PlacementManager; node names are their
    physical indices 0..n-1."""

    def __init__(self, nodes, slots):
        self.nodeStates = {c: NodeState(c, slots) for c in range(nodes)}
        self.jobStates = {}
        self.podNodeName = {}  # (job, worker index) -> node

    # placement_manager.go:337-411
    def release_slots(self, job_requests):
        for job in self.jobStates.values():
            num_workers = job_requests.get(job.name)
            if num_workers is None:
                # terminated: release all slots of the job (:344-360)
                for node_name, n_slots in job.nodeNumSlotsList:
                    node = self.nodeStates.get(node_name)
                    if node is not None:
                        node.freeSlots += n_slots
                        node.jobNumWorkers.pop(job.name, None)
                job.nodeNumSlotsList = []
            elif num_workers < job.numWorkers:
                # scaled down (:362-409): release from the last allocated node backwards
                to_release = job.numWorkers - num_workers
                while to_release > 0:
                    last = len(job.nodeNumSlotsList) - 1
                    entry = job.nodeNumSlotsList[last]
                    node = self.nodeStates.get(entry[0])
                    ok = node is not None
                    if entry[1] >= to_release:
                        entry[1] -= to_release
                        node.freeSlots += to_release
                        node.jobNumWorkers[job.name] = node.jobNumWorkers.get(job.name, 0) - to_release
                        to_release = 0
                    else:
                        to_release -= entry[1]
                        entry[1] = 0
                        if ok:
                            node.freeSlots += node.jobNumWorkers.get(job.name, 0)
                            node.jobNumWorkers[job.name] = 0
                    if job.nodeNumSlotsList[last][1] == 0:
                        job.nodeNumSlotsList.pop()
                    if ok and node.jobNumWorkers.get(job.name) == 0:
                        del node.jobNumWorkers[job.name]

    def previous_layout(self, job_index):
        """Post-release contents of the physical nodes as the placement round's input: per node
        (in physical order) the (round-local job index, workers) entries of the jobs in job_index."""
        out = []
        for c in sorted(self.nodeStates):
            out.append([(job_index[j], w) for j, w in self.nodeStates[c].jobNumWorkers.items() if j in job_index])
        return out

    # bindNodes :509-522 — logical node i takes the name of physical node phys_of_logical[i]
    def bind(self, layout, free_slots, phys_of_logical):
        """layout: (logical node, job name, workers) entries in bestFit's creation order."""
        n = len(self.nodeStates)
        total = {c: self.nodeStates[c].totalSlots for c in self.nodeStates}
        logical = []
        for i in range(n):
            nd = NodeState("TBD", total[i])
            nd.freeSlots = free_slots[i]
            logical.append(nd)
        for node, job, workers in layout:
            logical[node].jobNumWorkers[job] = workers
        new_states = {}
        for i in range(n):
            logical[i].name = phys_of_logical[i]
            new_states[logical[i].name] = logical[i]
        self.nodeStates = new_states

    # placement_manager.go:548-566
    def update_job_states(self):
        new = {}
        for c in sorted(self.nodeStates):
            node = self.nodeStates[c]
            for job_name, workers in node.jobNumWorkers.items():
                if job_name not in new:
                    new[job_name] = JobState(job_name)
                new[job_name].nodeNumSlotsList.append([node.name, workers])
                new[job_name].numWorkers += workers
        self.jobStates = new

    # placement_manager.go:571-617; returns (deletedWorkers, deletedLaunchers)
    def update_pod_node_name(self):
        new_table = {}
        deleted_workers = deleted_launchers = 0
        for job in self.jobStates.values():
            idx = 0
            deleted = 0
            for node_name, n_slots in job.nodeNumSlotsList:
                for _ in range(n_slots):
                    pod = (job.name, idx)
                    old = self.podNodeName.get(pod)
                    if old is not None and node_name != old:
                        deleted += 1
                        deleted_workers += 1
                    new_table[pod] = node_name
                    idx += 1
            if deleted == job.numWorkers:
                deleted_launchers += 1
        self.podNodeName = new_table
        return deleted_workers, deleted_launchers


def pod_table(manager, names):
    """pm.podNodeName as the run-length arrays voda_place takes (pods_off / pods_node / pods_slots):
    for every job of `names`, in order, the node of its worker pods 0, 1, 2, ... run-length coded."""
    off, node, slots = [0], [], []
    for job in names:
        idx = 0
        while (job, idx) in manager.podNodeName:
            nd = manager.podNodeName[(job, idx)]
            if len(node) > off[-1] and node[-1] == nd:
                slots[-1] += 1
            else:
                node.append(nd)
                slots.append(1)
            idx += 1
        off.append(len(node))
    return off, node, slots
