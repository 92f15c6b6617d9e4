"""CPU oracle for the gnn_b200 hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

Everything under ``oracle/`` exists to check the CUDA path, never to compute a
result that is shipped or measured as the product:

* ``pyg_shim``   restates the handful of torch_geometric 2.1.0 pieces the
                 reference imports (the reference pins torch-geometric==2.1.0 in
                 requirements.txt:2; the wheel is not in this image).
* ``restate``    numpy / torch-CPU restatement of the reference's hot path
                 (gnn/loaders/load.py, gnn/loaders/util.py, the GCN factories of
                 gnn/networks/networks.py, the epoch body of gnn/trainers/*.py).
* ``synth``      stateless synthetic genotype generator.
* ``ref_import`` imports the UNMODIFIED reference from /root/reference on top of
                 the shim (only possible in the build container).
* ``make_golden`` runs the reference itself and freezes tests/golden/*.npz, and
                 checks that ``restate`` reproduces the reference bit for bit.

Parity status: the reference ships no tests, golden vectors or known-answer
files for this path, so the oracle is pinned against
outputs of the reference's own Python run in the build container
(``make_golden.py``; fixtures committed under tests/golden/).  One deliberate,
documented difference: neighbour ties are ordered by (d^2, index) because
sklearn's own order depends on its thread count.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.
"""
