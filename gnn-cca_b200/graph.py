"""Cross-camera graph + edge-attribute construction on the device.

The reference has no function for this: the code sits inline in its epoch loops (train.py:257-259,
286-361,450; repeated at train.py:582-692 and inference.py:189-279).  ``build_graph_batch`` produces what
those lines hand to the MPN -- a PyG-``Batch``-like object with ``x``, ``edge_index`` (int64, bit-exact),
``edge_attr``, ``edge_labels``, ``y``, ``batch``, ``ptr`` -- from flat per-detection device tensors.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from .plan import _stream


class Batch:
    """The attributes of ``torch_geometric.data.Batch`` that the reference reads
    (train.py:454-458, inference.py:283-299)."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    # ``x``: the (column-normalised) node embeddings the reference hands to the MPN (train.py:257-259, 357).  A batch from
    # ``build_graph_batch`` keeps the caller's RAW embeddings and the 2048 column norms instead (``x_raw``, ``x_norm``); ``x`` is
    # materialised -- by the same kernel, bit for bit -- the first time somebody reads it.  ``MOTMPNet.forward`` does not: it feeds
    # ``x_raw`` to the node encoder and divides the weight columns of its first Linear by the norms (1 MB instead of a 200 MB rewrite).
    @property
    def x(self) -> torch.Tensor:
        xv = self.__dict__.get("_x")
        if xv is None:
            raw, norm = self.__dict__.get("x_raw"), self.__dict__.get("x_norm")
            if raw is None:
                raise AttributeError("x")
            lib = _lib.load()
            xv = torch.empty_like(raw)
            _lib.check(lib.gnncca_colnorm_apply(_lib.ptr(raw), _lib.ptr(xv), raw.size(0), raw.size(1), _lib.ptr(norm), _stream()), "colnorm_apply")
            self.__dict__["_x"] = xv
        return xv

    @x.setter
    def x(self, value) -> None:
        self.__dict__["_x"] = value
        self.__dict__.pop("x_raw", None)
        self.__dict__.pop("x_norm", None)

    def raw_x(self):
        """(x_raw, x_norm) when the normalised ``x`` has not been materialised or replaced, else None."""
        d = self.__dict__
        if d.get("_x") is None and d.get("x_raw") is not None:
            return d["x_raw"], d["x_norm"]
        return None

    @property
    def batch(self) -> torch.Tensor:
        """PyG's ``Batch.batch`` (frame index of every node), built from ``ptr`` on first use: the MPN never reads it."""
        b = self.__dict__.get("_batch_vec")
        if b is None:
            counts = self.ptr[1:] - self.ptr[:-1]
            b = torch.repeat_interleave(torch.arange(counts.numel(), device=self.ptr.device), counts, output_size=self.num_nodes)
            self.__dict__["_batch_vec"] = b
        return b

    @batch.setter
    def batch(self, value) -> None:
        self.__dict__["_batch_vec"] = value

    @property
    def num_nodes(self) -> int:
        d = self.__dict__
        return int((d["_x"] if d.get("_x") is not None else d["x_raw"]).size(0))

    @property
    def num_edges(self) -> int:
        return int(self.edge_index.size(1))

    def to(self, device):
        kw = {("x" if k == "_x" else k): (v.to(device) if torch.is_tensor(v) else v) for k, v in self.__dict__.items() if not k.startswith("_gnncca")}
        if kw.get("x") is not None:
            kw.pop("x_raw", None); kw.pop("x_norm", None)
        return Batch(**kw)


def column_norms(emb: torch.Tensor) -> torch.Tensor:
    """max(||emb[:, c]||_2, 1e-12) per column: the denominators of ``F.normalize(emb, p=2, dim=0)`` (train.py:257-259)."""
    lib = _lib.load()
    if not emb.is_cuda or emb.dtype != torch.float32 or emb.dim() != 2 or not emb.is_contiguous():
        raise ValueError("column_norms expects a contiguous 2-D float32 CUDA tensor")
    rows, cols = emb.shape
    scratch = torch.empty(int(lib.gnncca_colnorm_scratch_bytes(rows, cols)) // 4 + 1, dtype=torch.float32, device=emb.device)
    norm = torch.empty(cols, dtype=torch.float32, device=emb.device)
    _lib.check(lib.gnncca_colnorm_norms(_lib.ptr(emb), rows, cols, _lib.ptr(scratch), _lib.ptr(norm), _stream()), "colnorm_norms")
    return norm


def column_normalize(emb: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``F.normalize(emb, p=2, dim=0)`` (train.py:257-259): each feature COLUMN is divided by its L2 norm over
    all detections of the batch (clamped at 1e-12)."""
    lib = _lib.load()
    if not emb.is_cuda or emb.dtype != torch.float32 or emb.dim() != 2:
        raise ValueError("column_normalize expects a 2-D float32 CUDA tensor")
    emb = emb.contiguous()
    out = torch.empty_like(emb) if out is None else out
    rows, cols = emb.shape
    scratch = torch.empty(int(lib.gnncca_colnorm_scratch_bytes(rows, cols)) // 4 + 1, dtype=torch.float32, device=emb.device)
    _lib.check(lib.gnncca_colnorm(_lib.ptr(emb), _lib.ptr(out), rows, cols, _lib.ptr(scratch), _stream()), "colnorm")
    return out


_TOTALS = {}          # device index -> (pinned int64 [slots, 8], its numpy view, [next slot])
_TOTALS_SLOTS = 8


def _pinned_totals(dev: torch.device):
    """Eight int64 words of pinned (device-writable) host memory: the scan kernel stores the batch totals there itself."""
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    ent = _TOTALS.get(key)
    if ent is None:
        t = torch.zeros((_TOTALS_SLOTS, 8), dtype=torch.int64).pin_memory()
        ent = _TOTALS[key] = (t, t.numpy(), [0])
    t, arr, nxt = ent
    k = nxt[0]
    nxt[0] = (k + 1) % _TOTALS_SLOTS
    arr[k, 5] = 0
    arr[k, 6] = 0
    return t[k], arr[k]


def _poll_totals(arr, what: str = "graph_prepare_all"):
    """Spin until both scan blocks have set their check word (they store the values first and fence).  A kernel that never
    gets there (a launch failure upstream) ends the wait through the stream's own error after a bounded time."""
    import time
    magic = _lib.TOTALS_MAGIC
    spins, t0 = 0, None
    while arr[5] != magic or arr[6] != magic:
        spins += 1
        if spins & 0x3ff == 0:
            now = time.perf_counter()
            t0 = now if t0 is None else t0
            if now - t0 > 20.0:
                torch.cuda.current_stream().synchronize()           # raises the CUDA error, if that is what happened
                if arr[5] == magic and arr[6] == magic:
                    break
                raise _lib.GnnccaError("%s: the batch totals never arrived in host memory" % what)
    return tuple(int(v) for v in arr[:5])


_E_GUESS = {}         # (B, N) -> E of the last batch of that shape


def _alloc_outputs(E: int, width: int, dev):
    return (torch.empty((2, E), dtype=torch.int64, device=dev), torch.empty((E, width), dtype=torch.float32, device=dev),
            torch.empty(E, dtype=torch.float32, device=dev))


_SIZES = {}           # (B, N, R, F) -> byte sizes of the parts of the work buffer


def _as(t: torch.Tensor, dev: torch.device, dtype) -> torch.Tensor:
    if t.dtype is dtype and t.device == dev and t.is_contiguous():
        return t
    return t.to(device=dev, dtype=dtype).contiguous()


def build_graph_batch(graph_ptr: torch.Tensor, id_cam: torch.Tensor, ids: torch.Tensor, xw: torch.Tensor, yw: torch.Tensor,
                      max_dist: torch.Tensor, reid_embeds: torch.Tensor, node_embeds: torch.Tensor, mode: str = "full",
                      l2norm: bool = True, lazy_x: bool = True) -> Batch:
    """graph_ptr [B+1] int64 (detections of frame g are rows graph_ptr[g]:graph_ptr[g+1], in the row order of
    the frame's table), id_cam [N] int, ids [N] int, xw/yw [N] float64 ground-plane coordinates, max_dist [B]
    float64 (CONV_TO_M of the frame's dataset), reid_embeds [N,R] / node_embeds [N,F] float32 CNN outputs.
    ``mode``: 'full' | 'only_appearance' | 'only_dist' (train.py:335-344); ``l2norm`` = CNN_MODEL.L2norm.
    ``lazy_x``: keep ``node_embeds`` as given plus its column norms and materialise the normalised ``Batch.x`` only when it is
    read (see ``Batch.x``); False writes it here.

    Launch structure with ``l2norm`` (the default): ``gnncca_graph_prepare_all`` (frame tables, both scans and the ReID column norms
    in two launches; the node-embedding norms forked onto a side stream), the batch totals polled from pinned host memory the
    scan kernel writes itself, ``gnncca_build_graphs`` (which normalises the ReID rows while it stages them), join."""
    lib = _lib.load()
    dev = node_embeds.device
    if dev.type != "cuda":
        raise RuntimeError("gnn-cca_b200 builds graphs on CUDA tensors only (no CPU path)")
    attr_mode = _lib.ATTR_MODES[mode]
    graph_ptr = _as(graph_ptr, dev, torch.int64)
    B, N = int(graph_ptr.numel()) - 1, int(node_embeds.size(0))
    id_cam32 = _as(id_cam, dev, torch.int32)
    ids64 = _as(ids, dev, torch.int64)
    xw = _as(xw, dev, torch.float64)
    yw = _as(yw, dev, torch.float64)
    max_dist = _as(max_dist, dev, torch.float64)
    for name, t in (("reid_embeds", reid_embeds), ("node_embeds", node_embeds)):
        if t.dim() != 2 or t.dtype != torch.float32 or t.device != dev:
            raise ValueError("%s must be a 2-D float32 tensor on %s" % (name, dev))
    R, F = int(reid_embeds.size(1)), int(node_embeds.size(1))
    # one work buffer: [frame tables | reciprocal ReID norms | ReID partial sums | node partial sums], 256-byte aligned parts
    sizes = _SIZES.get((B, N, R, F))
    if sizes is None:
        up = lambda n: (int(n) + 255) & ~255
        sizes = _SIZES[(B, N, R, F)] = (up(lib.gnncca_graph_prepare_scratch_bytes(B, N)), up(4 * R),
                                        up(lib.gnncca_colnorm_scratch_bytes(N, R)), up(lib.gnncca_colnorm_scratch_bytes(N, F)))
        if len(_SIZES) > 64:
            _SIZES.clear()
    nbytes = sizes[0]
    work = torch.empty(sum(sizes) // 8, dtype=torch.int64, device=dev)
    scratch_p = work.data_ptr()
    edge_ptr = torch.empty(B + 1, dtype=torch.int64, device=dev)
    stream = _stream()
    need_reid = mode != "only_dist"
    x_kw, rnorm_p, join, pre_out, guess = {}, None, False, None, None

    # shapes the single prepare call takes (others: the separate launches below, same results)
    fused_ok = N <= 65535 * 128 and (not need_reid or (R % 4 == 0 and R <= 1024 and reid_embeds.data_ptr() % 16 == 0))
    if l2norm and B > 0 and N > 0 and fused_ok:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("build_graph_batch cannot be captured into a CUDA graph: the sizes of its outputs come from the device")
        reid = reid_embeds.contiguous()
        node = node_embeds.contiguous()
        rnorm_p = scratch_p + sizes[0] if need_reid else None
        x_norm = torch.empty(F, dtype=torch.float32, device=dev) if lazy_x else None
        tot_t, tot = _pinned_totals(dev)
        _lib.check(lib.gnncca_graph_prepare_all(
            graph_ptr.data_ptr(), id_cam32.data_ptr(), B, N, edge_ptr.data_ptr(), scratch_p, nbytes,
            node.data_ptr() if lazy_x else None, F, _lib.ptr(x_norm), (scratch_p + sizes[0] + sizes[1] + sizes[2]) if lazy_x else None,
            reid.data_ptr() if need_reid else None, R, rnorm_p, (scratch_p + sizes[0] + sizes[1]) if need_reid else None,
            tot_t.data_ptr(), stream), "graph_prepare_all")
        # the outputs are sized by E, which is on its way from the device: allocate for the E the last batch of this shape had
        # while waiting (a loader's batches repeat their shape); a wrong guess costs three cached allocations
        guess = _E_GUESS.get((B, N))
        if guess is not None:
            pre_out = _alloc_outputs(guess, 4 if mode == "full" else 2, dev)
        join = lazy_x
        if lazy_x:
            x_kw = dict(x_raw=node, x_norm=x_norm)
        else:
            x_kw = dict(x=column_normalize(node))            # enqueued while the totals are on their way
        n_tiles, E, max_n, max_e, cam_unsorted = _poll_totals(tot)
    else:
        if l2norm:
            reid = column_normalize(reid_embeds)
            x_kw = dict(x_raw=node_embeds.contiguous(), x_norm=column_norms(node_embeds.contiguous())) if lazy_x \
                else dict(x=column_normalize(node_embeds))
        else:
            reid, x_kw = reid_embeds.contiguous(), dict(x=node_embeds.contiguous())
        _lib.check(lib.gnncca_graph_edge_counts(_lib.ptr(graph_ptr), _lib.ptr(id_cam32), B, N, _lib.ptr(edge_ptr),
                                                scratch_p, nbytes, stream), "graph_edge_counts")
        # the output size is data dependent: one 40-byte read back (the reference syncs several times per FRAME)
        if B > 0:
            off = (lib.gnncca_graph_tile_total_ptr(scratch_p, B, N) - scratch_p) // 8
            n_tiles, E, max_n, max_e, cam_unsorted = (int(v) for v in work[off:off + 5].tolist())
        else:
            n_tiles = E = max_n = max_e = cam_unsorted = 0
    caps = (max_n, max_e)      # exact frame maxima: sizes the fused all-steps kernel without guessing

    if pre_out is not None and guess == E:
        edge_index, edge_attr, edge_labels = pre_out
    else:
        edge_index, edge_attr, edge_labels = _alloc_outputs(E, 4 if mode == "full" else 2, dev)
        if len(_E_GUESS) > 64:
            _E_GUESS.clear()
        _E_GUESS[(B, N)] = E
    _lib.check(lib.gnncca_build_graphs(graph_ptr.data_ptr(), edge_ptr.data_ptr(), B, N, E, n_tiles, id_cam32.data_ptr(),
                                       ids64.data_ptr(), xw.data_ptr(), yw.data_ptr(), max_dist.data_ptr(), reid.data_ptr(),
                                       rnorm_p, R, attr_mode, max_n, scratch_p, edge_index.data_ptr(),
                                       edge_attr.data_ptr(), edge_labels.data_ptr(), stream), "build_graphs")
    if join:
        _lib.check(lib.gnncca_graph_join(stream), "graph_join")
    # "trusted" = source-major by construction: true only when every frame's rows are in ascending camera order (the reference
    # enumerates camera by camera, train.py:292-299, so interleaved cameras give an edge_index[0] that is NOT non-decreasing);
    # otherwise the forward takes its checked path (order flags read back, stable sort + permutation)
    return Batch(edge_index=edge_index, edge_attr=edge_attr, edge_labels=edge_labels, y=ids64, **x_kw,
                 ptr=graph_ptr, edge_ptr=edge_ptr, num_graphs=B, gnncca_caps=caps, gnncca_trusted=cam_unsorted == 0)
