"""The C-ABI library loads on a machine without a GPU and exports every symbol include/gnncca.h declares
(no compute calls here: those are the -m gpu tests)."""
import ctypes
import importlib
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "gnncca.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gnncca_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_hot_path_entry_points():
    names = declared_functions()
    for must in ("gnncca_csr_from_edge_index", "gnncca_node_encoder", "gnncca_mp_step_fwd", "gnncca_mp_fused_fwd", "gnncca_build_graphs",
                 "gnncca_colnorm", "gnncca_edge_bwd", "gnncca_node_agg_bwd", "gnncca_finalize_grads"):
        assert must in names
    assert len(names) >= 30


def test_library_exports_every_declared_symbol_and_bindings_cover_them():
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    lib = lib_mod.load()
    raw = ctypes.CDLL(lib_mod.LIB_PATH)
    for name in declared_functions():
        assert hasattr(raw, name), "libgnncca.so does not export %s" % name
        assert name in lib_mod.SIGNATURES, "gnn-cca_b200/_lib.py has no ctypes signature for %s" % name
    assert set(lib_mod.SIGNATURES) <= set(declared_functions())
    assert lib.gnncca_abi_version() == 1
    assert lib.gnncca_status_string(-2).decode().startswith("unsupported")
    assert lib.gnncca_train_blocks(1000, lib_mod.BLOCKS_PER_EDGE) == 4     # pure host arithmetic


def test_struct_layouts_match_the_header():
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    assert ctypes.sizeof(lib_mod.MpCfg) == 40
    assert ctypes.sizeof(lib_mod.MpWeights) == 12 * 8 + 8
    assert ctypes.sizeof(lib_mod.FinSeg) == 2 * 8 + 8 * 4


def test_only_sm100a_code_is_built():
    """No multi-arch / PTX-JIT fallback: the library holds sm_100a SASS only."""
    import shutil
    import subprocess
    lib_mod = importlib.import_module("gnn-cca_b200._lib")
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        return
    out = subprocess.run([cuobjdump, "-lelf", lib_mod.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
