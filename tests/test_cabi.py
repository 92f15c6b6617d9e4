"""The C-ABI library: loads without a GPU, exports every symbol include/gnnb.h declares, and refuses to
compute without a B200 (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

from gnn_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gnnb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gnnb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 25
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for name in names:
        assert hasattr(handle, name), f"{name} declared in gnnb.h but not exported"
    assert sorted(_lib.PROTOTYPES) == names, "gnn_b200/_lib.py prototypes out of step with include/gnnb.h"


def test_version_and_error_string():
    lib = _lib.lib()
    assert lib.gnnb_version() >= 100
    assert isinstance(lib.gnnb_last_error(), bytes)


def test_workspace_queries_need_no_gpu():
    assert _lib.workspace_bytes("gnnb_graph_workspace_bytes", 1000, 8) > 2 * 1000 * 8 * 16
    assert _lib.workspace_bytes("gnnb_csr_workspace_bytes", 1000, 9000) > 2 * (9000 + 1000) * 16
    assert _lib.workspace_bytes("gnnb_knn_workspace_bytes", 1000, 1000, 128, 32, 0) > 1000 * 32 * 12
    assert _lib.workspace_bytes("gnnb_colsum_workspace_bytes", 100) == 1024 * 100 * 4
    with pytest.raises(_lib.GnnbError):
        _lib.workspace_bytes("gnnb_knn_workspace_bytes", 10, 10, 10, 65, 0)      # k above GNNB_KNN_MAX_K


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_compute_fails_loudly_without_gpu():
    assert _lib.lib().gnnb_check_device() < 0
    with pytest.raises(_lib.GnnbError, match="no CPU fallback|no CUDA device|not sm_100"):
        _lib.call("gnnb_row_sqnorm_i8", None, 0, 0, 1, None, None)
    from gnn_b200 import ops
    with pytest.raises(_lib.GnnbError):
        ops.knn(torch.zeros((4, 8), dtype=torch.int8), 2)
    from gnn_b200.networks.gcn_conv import GCNConv
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        GCNConv(8, 4)(torch.zeros(4, 8), torch.zeros((2, 0), dtype=torch.long))


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under gnn_b200/ may import or execute it."""
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gnn_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                if re.search(r"^\s*(from|import)\s+oracle\b|from\s+\.\.?oracle|/root/reference", src, flags=re.M):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad
