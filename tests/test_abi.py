"""C-ABI boundary (CPU, no compute): the header parses, liblgk.so loads and exports every declared symbol, the
argument validation paths answer without touching a device, and the product refuses CPU tensors."""
import ctypes
import re
import subprocess

import pytest
import torch

from lesion_gnn_b200 import _lib


def test_header_declares_expected_groups():
    protos = _lib.parse_header()
    for name in ("lgk_csr_build", "lgk_batch_ptr", "lgk_knn_graph", "lgk_spmm", "lgk_spmm_max_bwd", "lgk_gcn_norm",
                 "lgk_gat_fwd", "lgk_gat_bwd_dst", "lgk_gat_bwd_src", "lgk_pool_fwd", "lgk_pool_bwd", "lgk_linear_fwd",
                 "lgk_linear_bwd_x", "lgk_linear_bwd_w", "lgk_adam_step", "lgk_last_error"):
        assert name in protos
    # no torch / C++ types at the boundary: only pointers and scalars
    text = re.sub(r"/\*.*?\*/", " ", open(_lib.HEADER).read(), flags=re.S)
    assert "std::" not in text and "at::" not in text and "Tensor" not in text


def test_every_declared_symbol_is_exported(liblgk):
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIBRARY], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(lgk_\w+)", out))
    declared = set(_lib.parse_header())
    assert declared <= exported, declared - exported
    assert exported <= declared, f"undeclared exports: {exported - declared}"


def test_library_is_sm100a_only(liblgk):
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.LIBRARY], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_bad_arguments_return_status_without_a_device(liblgk):
    L = liblgk
    assert L.query("lgk_version") >= 100
    rc = L._dll.lgk_spmm(None, None, None, None, None, None, None, 1.0, 0, None, None, None, None, 4, 4, 0, 0, None)
    assert rc == -1 and "null pointer" in L.last_error()
    rc = L._dll.lgk_spmm(None, None, None, None, None, None, None, 1.0, 7, ctypes.c_void_p(16), None, ctypes.c_void_p(32),
                         None, 4, 4, 0, 0, None)
    assert rc == -1
    rc = L._dll.lgk_knn_graph(ctypes.c_void_p(16), 0, 2, ctypes.c_void_p(16), ctypes.c_void_p(16), 1, 4, 100, 1,
                              ctypes.c_void_p(16), 4, None)
    assert rc == -2 and "k must be" in L.last_error()
    assert L.query("lgk_csr_workspace_bytes", 100, 1000) >= 16 * 1000
    with pytest.raises(RuntimeError, match="lgk_gat_fwd failed"):
        L.call("lgk_gat_fwd", None, None, None, None, None, None, None, None, None, 4, 2, 3, 0.2, 1, 0.0, 0, None, 0, None)


def test_product_refuses_cpu_tensors(liblgk):
    import lesion_gnn_b200 as lg
    x = torch.randn(4, 8)
    ei = torch.tensor([[0, 1], [1, 0]])
    with pytest.raises(RuntimeError, match="no CPU path"):
        lg.nn.GCNConv(8, 8)(x, ei)
    with pytest.raises(RuntimeError, match="no CPU path"):
        lg.nn.global_mean_pool(x, torch.zeros(4, dtype=torch.long))
    with pytest.raises(RuntimeError, match="no CPU path"):
        lg.transforms.knn_graph(torch.rand(4, 2), 2)


def test_product_never_imports_the_oracle():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "lesion_gnn_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
