"""CPU restatement of the reference's edge decisions, pruning and clustering (TEST INFRASTRUCTURE ONLY).

Only ``tests/`` and ``__graft_entry__.smoke()`` may import this module; the product path (``gnn-cca_b200/postprocess.py``
-> ``libgnncca.so``) never does.

Restates inference.py:286-312 with ``remove_edges_single_direction`` (/root/reference/libs/utils.py:387-404) and
``compute_SCC_and_Clusters`` (utils.py:295-317).  The latter calls networkx 2.x ``strongly_connected_components``
(third-party; published algorithm: Tarjan / Nuutila); here an iterative Tarjan in plain Python.  Cluster NUMBERING in
the reference follows networkx's traversal order (components sorted by size, isolated nodes appended); it is consumed
only by permutation-invariant metrics, so parity is on the PARTITION: ``canonical`` renumbers any labelling densely in
order of each cluster's smallest node.  Pinned by ``tests/golden/post_*.npz`` (reference functions run unmodified with
networkx, ``oracle/gen_golden.py`` ``gen_post``).
"""
from __future__ import annotations

import numpy as np


def decisions(logits: np.ndarray) -> np.ndarray:
    """(sigmoid(x) >= 0.5) * 1 in float32 (inference.py:289-291)."""
    x = np.asarray(logits, np.float32).reshape(-1)
    with np.errstate(over="ignore"):
        return ((np.float32(1) / (np.float32(1) + np.exp(-x, dtype=np.float32))) >= np.float32(0.5)).astype(np.int64)


def remove_edges_single_direction(predictions: np.ndarray, edge_list: np.ndarray) -> np.ndarray:
    """utils.py:387-404: an active edge (u, v) whose reverse (v, u) is not active is switched off."""
    pred = np.asarray(predictions, np.int64).copy()
    active = {(int(edge_list[0][e]), int(edge_list[1][e])) for e in np.nonzero(pred == 1)[0]}
    for e in np.nonzero(pred == 1)[0]:
        if (int(edge_list[1][e]), int(edge_list[0][e])) not in active:
            pred[e] = 0
    return pred


def strongly_connected_components(n_nodes: int, edges):
    """Tarjan (iterative).  Returns a list of sorted node lists; every node 0..n_nodes-1 appears once (utils.py:299-308
    appends the nodes networkx never saw as singletons)."""
    adj = [[] for _ in range(n_nodes)]
    for u, v in edges:
        adj[u].append(v)
    index = [-1] * n_nodes
    low = [0] * n_nodes
    on = [False] * n_nodes
    stack, comps, counter = [], [], 0
    for root in range(n_nodes):
        if index[root] != -1:
            continue
        work = [(root, 0)]
        while work:
            v, i = work.pop()
            if i == 0:
                index[v] = low[v] = counter
                counter += 1
                stack.append(v)
                on[v] = True
            recurse = False
            for j in range(i, len(adj[v])):
                w = adj[v][j]
                if index[w] == -1:
                    work.append((v, j + 1))
                    work.append((w, 0))
                    recurse = True
                    break
                if on[w]:
                    low[v] = min(low[v], index[w])
            if recurse:
                continue
            if low[v] == index[v]:
                comp = []
                while True:
                    w = stack.pop()
                    on[w] = False
                    comp.append(w)
                    if w == v:
                        break
                comps.append(sorted(comp))
            if work:
                low[work[-1][0]] = min(low[work[-1][0]], low[v])
    return comps


def canonical(ids: np.ndarray) -> np.ndarray:
    """Dense renumbering of a labelling in order of each cluster's smallest node (partition-preserving)."""
    ids = np.asarray(ids).reshape(-1)
    out = np.empty(ids.shape[0], np.int64)
    seen = {}
    for i, c in enumerate(ids.tolist()):
        if c not in seen:
            seen[c] = len(seen)
        out[i] = seen[c]
    return out


def clusters(n_nodes: int, predictions: np.ndarray, edge_list: np.ndarray):
    """compute_SCC_and_Clusters(nx.DiGraph(active edges), n_nodes) -> (canonical ids [n_nodes], n_clusters)."""
    act = np.nonzero(np.asarray(predictions) == 1)[0]
    comps = strongly_connected_components(n_nodes, [(int(edge_list[0][e]), int(edge_list[1][e])) for e in act])
    ids = np.empty(n_nodes, np.int64)
    for c, comp in enumerate(comps):
        ids[comp] = c
    return canonical(ids), len(comps)
