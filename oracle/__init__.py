"""CPU oracle for the lesion-gnn message-passing hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``lesion_gnn_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and only as the checker / the
timed CPU baseline, never as the product path.

PARITY UNPINNED: the arithmetic of this path lives in third-party wheels that
are not vendored under /root/reference and are not installable here
(torch-geometric==2.5.1, torch-scatter==2.1.2, torch-sparse==0.6.18,
torch-cluster==1.6.3; /root/reference/requirements.lock:405-414), and the
reference's own tests hold no golden vector for any layer on the path
(/root/reference/test/test_transforms.py pins only GaussianDistance, which is
reproduced in tests/test_oracle.py).  The oracle therefore restates the
published PyG 2.5.1 algorithms op for op (gather/scatter formulation) and is
cross-checked against an independent dense-matrix formulation
(oracle/dense.py), torch.autograd.gradcheck in fp64, and scipy's cKDTree for
the kNN graph.  If ``import torch_geometric`` ever succeeds on a box,
tests/test_oracle.py::test_against_real_pyg asserts oracle == PyG.
"""
