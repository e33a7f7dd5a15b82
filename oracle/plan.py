"""Message-free restatement of the L2GMap halo plan (TEST INFRASTRUCTURE).

spmv/L2GMap.cpp:346-479 builds the plan with MPI_Allgather, MPI_Alltoall and
MPI_Neighbor_alltoallv.  Here the same artefacts are derived from
GLOBAL knowledge -- every rank's local size and sorted ghost list at once -- with no
transport at all, so this file shares no code (and no possible ordering bug) with the
thread-rank MPI shim that both the compiled reference (oracle/_ref, tier 2) and the
product's host layer link.  tests/test_oracle.py compares the three.

Also restates the data movement of L2GMap::update (L2GMap.cpp:502-526) and
L2GMap::reverse_update (collective variant, L2GMap.cpp:908-923) on plain numpy vectors.
"""
import bisect

import numpy as np


def plans(local_sizes, ghost_lists):
    """-> list (one per rank) of dicts with the keys of tests/cases.py::PLAN_KEYS.

    local_sizes[r]: owned entries of rank r; ghost_lists[r]: its ghosts, global, ascending."""
    size = len(local_sizes)
    ranges = [0]
    for n in local_sizes:                                   # L2GMap.cpp:351-356
        ranges.append(ranges[-1] + int(n))
    owner_of = lambda g: bisect.bisect_right(ranges, g) - 1  # noqa: E731  (:374 upper_bound - 1)
    # wants[r][q] = the ghosts of rank r that rank q owns, in r's (ascending) ghost order
    wants = []
    for r in range(size):
        gl = [int(g) for g in ghost_lists[r]]
        if any(b < a for a, b in zip(gl, gl[1:])):
            raise ValueError("Ghosts must be sorted")       # :362-363
        per = [[] for _ in range(size)]
        for g in gl:
            if ranges[r] <= g < ranges[r + 1]:
                raise ValueError("Ghost index in local range")  # :371-372
            per[owner_of(g)].append(g)
        wants.append(per)
    out = []
    for r in range(size):
        nbrs, send_count, recv_count, index = [], [], [], []
        for q in range(size):                               # :390-412, ascending rank
            mine_from_q = len(wants[r][q])                  # ghost_count[q]
            theirs_from_me = len(wants[q][r])               # remote_count[q] (the Alltoall, :386-388)
            if mine_from_q > 0 or theirs_from_me > 0:
                nbrs.append(q)
                send_count.append(mine_from_q)
                recv_count.append(theirs_from_me)
                index.extend(g - ranges[r] for g in wants[q][r])  # :441-447: owner-local, requester's order
        if not nbrs:
            send_count, recv_count = [0], [0]               # :421-425
        send_off, recv_off = [0], [0]
        for c in send_count:                                # :432-438
            send_off.append(send_off[-1] + c)
        for c in recv_count:
            recv_off.append(recv_off[-1] + c)
        send_off = [s + int(local_sizes[r]) for s in send_off]  # :460-461
        out.append({
            "ranges": np.array(ranges, np.int64), "ghosts": np.array(ghost_lists[r], np.int64).reshape(-1),
            "neighbours": np.array(nbrs, np.int32), "send_count": np.array(send_count, np.int32),
            "recv_count": np.array(recv_count, np.int32), "send_offset": np.array(send_off, np.int32),
            "recv_offset": np.array(recv_off, np.int32), "indexbuf": np.array(index, np.int32),
            "local_size": int(local_sizes[r]), "num_ghosts": len(ghost_lists[r]), "global_size": ranges[-1],
            "global_offset": ranges[r],
        })
    return out


def update(plan_list, vectors):
    """Forward halo (L2GMap.cpp:502-526): every rank packs vec[indexbuf] and neighbour i's block lands in
    that neighbour's ghost tail.  vectors[r]: numpy array of local_size + num_ghosts entries, updated in place."""
    packed = [v[p["indexbuf"]] if len(p["indexbuf"]) else np.zeros(0, v.dtype) for p, v in zip(plan_list, vectors)]
    for r, p in enumerate(plan_list):
        for i, q in enumerate(p["neighbours"]):
            pq = plan_list[q]
            j = list(pq["neighbours"]).index(r)             # I am neighbour j of q
            blk = packed[q][pq["recv_offset"][j]: pq["recv_offset"][j + 1]]   # what q sends to me
            assert len(blk) == p["send_count"][i]
            vectors[r][p["send_offset"][i]: p["send_offset"][i] + p["send_count"][i]] = blk


def reverse_update(plan_list, vectors):
    """Reverse halo (L2GMap.cpp:908-923): every rank sends its ghost tail back to the owners, the owner
    accumulates `vec[indexbuf[k]] += received[k]` for k ascending (:920-921).  In place."""
    tails = [v[p["local_size"]:].copy() for p, v in zip(plan_list, vectors)]
    for r, p in enumerate(plan_list):
        recv = np.zeros(len(p["indexbuf"]), vectors[r].dtype)
        for i, q in enumerate(p["neighbours"]):
            pq = plan_list[q]
            j = list(pq["neighbours"]).index(r)
            a = pq["send_offset"][j] - pq["local_size"]     # q's ghosts owned by me, in q's tail
            blk = tails[q][a: a + pq["send_count"][j]]
            assert len(blk) == p["recv_count"][i]
            recv[p["recv_offset"][i]: p["recv_offset"][i + 1]] = blk
        for k, idx in enumerate(p["indexbuf"]):
            vectors[r][idx] += recv[k]
