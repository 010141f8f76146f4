"""Source-sorted int32 CSR of a batched graph, built on the device once per forward
(replaces the implicit ``x[row]`` / ``x[col]`` / ``scatter(..., row)`` index handling of models/mpn.py:44-99)."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib


def _stream() -> int:
    """Raw handle of torch's current CUDA stream (the fast path: no Stream object is built)."""
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


class GraphPlan:
    """rowptr[N+1], src[E], dst[E] (int32, source-sorted) and, when the caller's edge order is not
    source-major, ``perm`` (sorted position -> original edge id)."""

    def __init__(self, edge_index: torch.Tensor, num_nodes: int):
        if edge_index.dim() != 2 or edge_index.size(0) != 2 or edge_index.dtype != torch.int64:
            raise ValueError("edge_index must be an int64 tensor of shape [2, E]")
        if not edge_index.is_cuda:
            raise RuntimeError("gnn-cca_b200 runs on CUDA tensors only (no CPU path)")
        self.lib = _lib.load()
        self.edge_index = edge_index.contiguous()
        self.N, self.E = int(num_nodes), int(edge_index.size(1))
        dev = edge_index.device
        self.src = torch.empty(self.E, dtype=torch.int32, device=dev)
        self.dst = torch.empty(self.E, dtype=torch.int32, device=dev)
        self.rowptr = torch.empty(self.N + 1, dtype=torch.int32, device=dev)
        self.flags = torch.empty(_lib.CSR_NFLAGS, dtype=torch.int32, device=dev)
        self.perm: Optional[torch.Tensor] = None
        self.raw_edge_index = None      # CSR form of gnncca_mp_fused_fwd
        self.edge_ptr = None
        self._flags_host = None     # pinned read-back buffer + event, created on first use only
        self._event = None
        self._resolved = False
        self.last_path = None
        _lib.check(self.lib.gnncca_csr_from_edge_index(_lib.ptr(self.edge_index), self.N, self.E, _lib.ptr(self.src),
                                                       _lib.ptr(self.dst), _lib.ptr(self.rowptr), _lib.ptr(self.flags),
                                                       _stream()), "csr_from_edge_index")
        self._requested = False

    @property
    def resolved(self) -> bool:
        return self._resolved

    def request_flags(self) -> None:
        """Enqueue the 16-byte read back of the flags behind everything launched so far (non-blocking)."""
        if self._flags_host is None:
            self._flags_host = torch.empty(_lib.CSR_NFLAGS, dtype=torch.int32, pin_memory=True)
            self._event = torch.cuda.Event()
        self._flags_host.copy_(self.flags, non_blocking=True)
        self._event.record()
        self._requested = True

    def check_deferred(self) -> None:
        """For plans used WITHOUT waiting for their flags (batches trusted by construction; the step kernels take the flags as
        their guard and are no-ops on an invalid CSR): if a read back requested by an earlier forward has completed, look at
        it now and raise when the trust was misplaced.  Never blocks."""
        if not self._requested or self._event is None or not self._event.query():
            return
        self._requested = False
        if int(self._flags_host[1]):
            raise IndexError("edge_index has entries outside [0, num_nodes=%d) (found after a forward on a batch marked trusted: "
                             "its outputs are invalid)" % self.N)
        if int(self._flags_host[3]):
            raise RuntimeError("a batch marked gnncca_trusted is not symmetric (an edge without its reverse): the backward that used it "
                               "computed wrong gradients; clear data.gnncca_trusted so that the checked path sorts by target")
        if int(self._flags_host[0]):
            raise RuntimeError("a batch marked gnncca_trusted is not source-major: the forward that used it was a no-op and its "
                               "outputs are invalid; clear data.gnncca_trusted so that the checked path sorts the edges")

    def frames_overflowed(self) -> bool:
        """flags[2]: raised by gnncca_mp_fused_fwd when a frame did not fit its capacity class (valid after
        ``source_major()``)."""
        return bool(int(self._flags_host[2]))

    def key(self):
        """Identity of the edge_index this CSR was built from (tensor storage, shape and in-place version)."""
        t = self.edge_index
        return (t.data_ptr(), tuple(t.shape), t._version, self.N)

    def clear_overflow(self) -> None:
        self.flags[2:3].zero_()

    def source_major(self) -> bool:
        """Wait for the order / range flags (a 16-byte read back) and report whether the caller's edge order is
        already source-major (what train.py:292-299 produces).  Raises on out-of-range indices."""
        if not self._requested:
            self.request_flags()
        self._event.synchronize()
        self._requested = False
        if int(self._flags_host[1]):
            raise IndexError("edge_index has entries outside [0, num_nodes=%d)" % self.N)
        return not int(self._flags_host[0])

    def resolve(self) -> "GraphPlan":
        """Make the CSR valid: read the flags and run the stable sort when the edges are not source-major."""
        if self._resolved:
            return self
        unsorted = not self.source_major()
        if unsorted:
            dev = self.src.device
            self.perm = torch.empty(self.E, dtype=torch.int32, device=dev)
            nbytes = int(self.lib.gnncca_csr_sort_workspace_bytes(self.N, self.E))
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.check(self.lib.gnncca_csr_sort(self.N, self.E, _lib.ptr(self.src), _lib.ptr(self.dst), _lib.ptr(self.perm),
                                                _lib.ptr(self.rowptr), _lib.ptr(ws), nbytes, _stream()), "csr_sort")
            self.flags.zero_()        # the CSR is valid now: lift the guard the step kernels look at
        self._resolved = True
        return self

    # edge-row permutations between the caller's order and the sorted order
    def to_sorted(self, rows: torch.Tensor) -> torch.Tensor:
        if self.perm is None:
            return rows
        out = torch.empty_like(rows)
        _lib.check(self.lib.gnncca_permute_rows(_lib.ptr(rows), _lib.ptr(out), _lib.ptr(self.perm), rows.size(0),
                                                rows.size(1), 0, _stream()), "permute_rows")
        return out

    def to_caller(self, rows: torch.Tensor) -> torch.Tensor:
        if self.perm is None:
            return rows
        out = torch.empty_like(rows)
        _lib.check(self.lib.gnncca_permute_rows(_lib.ptr(rows), _lib.ptr(out), _lib.ptr(self.perm), rows.size(0),
                                                rows.size(1), 1, _stream()), "permute_rows")
        return out


_TRUSTED_FLAGS = {}


class TrustedPlan:
    """Stand-in for ``GraphPlan`` when the batch was built by ``build_graph_batch`` (source-major, frames closed,
    per-frame edge offsets known): the fused kernel reads the caller's int64 ``edge_index`` itself (RAW form of
    gnncca_mp_fused_fwd), no CSR arrays exist and nothing is read back."""

    resolved = True
    perm = None
    rowptr = src = dst = None

    def __init__(self, edge_index: torch.Tensor, num_nodes: int, edge_ptr: torch.Tensor):
        if edge_index.dim() != 2 or edge_index.size(0) != 2 or edge_index.dtype != torch.int64 or not edge_index.is_cuda:
            raise ValueError("edge_index must be a CUDA int64 tensor of shape [2, E]")
        dev = edge_index.device
        self.edge_index = self.raw_edge_index = edge_index.contiguous()
        self.edge_ptr = edge_ptr.to(device=dev, dtype=torch.int64).contiguous()
        self.N, self.E = int(num_nodes), int(edge_index.size(1))
        key = (dev.type, dev.index)
        if key not in _TRUSTED_FLAGS:          # write-only scratch for the kernel's overflow flag (never read back)
            _TRUSTED_FLAGS[key] = torch.zeros(_lib.CSR_NFLAGS, dtype=torch.int32, device=dev)
        self.flags = _TRUSTED_FLAGS[key]
        self._resolved = True
        self.last_path = None

    def key(self):
        t = self.edge_index
        return (t.data_ptr(), tuple(t.shape), t._version, self.N)

    def to_sorted(self, rows: torch.Tensor) -> torch.Tensor:
        return rows

    def to_caller(self, rows: torch.Tensor) -> torch.Tensor:
        return rows
