"""Import the UNMODIFIED reference modules from /root/reference (TEST INFRASTRUCTURE ONLY).

Only usable in the build container: ``/root/reference`` does not exist on the GPU box, so nothing that
runs there (``-m gpu`` tests, ``smoke()``, ``bench.py``) may call this at run time; they use the
restatements in ``oracle/mpn_oracle.py`` / ``oracle/graph_oracle.py`` and the committed
``tests/golden`` vectors instead.

The reference needs third-party modules that are absent from this image.  They are
replaced, for the reference's import only, by small stand-ins installed in ``sys.modules``:

* ``torch_scatter``  (2.0.8) -> ``oracle.mpn_oracle.scatter_add/mean/max``
* ``torch_geometric.data.Data / Batch`` (2.0.1) -> ``oracle.graph_oracle.Data / Batch``
* ``matplotlib(.pyplot)`` -> empty namespace (plots are never reached)
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "mpn.py"))


def _install_stubs() -> None:
    from . import mpn_oracle, graph_oracle

    if "torch_scatter" not in sys.modules:
        ts = types.ModuleType("torch_scatter")
        ts.scatter_add = mpn_oracle.scatter_add
        ts.scatter_mean = mpn_oracle.scatter_mean
        ts.scatter_max = mpn_oracle.scatter_max
        sys.modules["torch_scatter"] = ts
    if "torch_geometric" not in sys.modules:
        tg = types.ModuleType("torch_geometric")
        tgd = types.ModuleType("torch_geometric.data")
        tgd.Data = graph_oracle.Data
        tgd.Batch = graph_oracle.Batch
        tg.data = tgd
        sys.modules["torch_geometric"] = tg
        sys.modules["torch_geometric.data"] = tgd
    if "matplotlib" not in sys.modules:
        try:
            importlib.import_module("matplotlib")
        except Exception:
            mpl = types.ModuleType("matplotlib")
            plt = types.ModuleType("matplotlib.pyplot")
            mpl.pyplot = plt
            mpl.use = lambda *a, **k: None
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = plt


def import_reference(module: str):
    """Import e.g. 'models.mpn' or 'inference' from /root/reference, unmodified."""
    if not reference_available():
        raise RuntimeError("reference tree not present (expected on the GPU box); use the oracle restatement")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    # A drop-in shadow of ``models`` (gnn-cca_b200/dropin) may already be imported; never mix them up.
    for name in ("models", "models.mpn", "models.mlp"):
        mod = sys.modules.get(name)
        if mod is not None and not str(getattr(mod, "__file__", "")).startswith(REFERENCE_ROOT):
            del sys.modules[name]
    return importlib.import_module(module)


def reference_def(rel_path: str, name: str, namespace: dict):
    """Compile ONE top-level ``def``/``class`` of a reference file, unmodified, into ``namespace`` and return it.

    For modules whose import needs packages this image lacks (train.py: cv2, skimage, torchreid, networkx ...): the
    function's own source text is taken from the file with ``ast`` (exact line range) and executed as is."""
    import ast

    if not reference_available():
        raise RuntimeError("reference tree not present (expected on the GPU box); use the oracle restatement")
    path = os.path.join(REFERENCE_ROOT, rel_path)
    with open(path, "r") as f:
        text = f.read()
    for node in ast.parse(text).body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name == name:
            lines = text.splitlines()[node.lineno - 1:node.end_lineno]
            exec(compile("\n" * (node.lineno - 1) + "\n".join(lines), path, "exec"), namespace)
            return namespace[name]
    raise KeyError("%s not found in %s" % (name, rel_path))
