"""gnn_b200 -- sm_100a (B200) implementation of the message-passing hot path of ragsGo/GNN.

Module layout mirrors the reference's ``gnn`` package for the path it covers:

    gnn_b200.loaders.load      create_graph, PlainGraph, load_data          (gnn/loaders/load.py)
    gnn_b200.loaders.util      split_dataset_graph, split_dataset, naive_partition   (gnn/loaders/util.py)
    gnn_b200.networks.*        GCNConv, Sequential, create_network_* , Ensemble      (gnn/networks/*.py)
    gnn_b200.trainers.*        train, train_batches, train_ensemble2, get_args_tuple (gnn/trainers/*.py)
    gnn_b200.ops               tensor-level wrappers over the C ABI of libgnn_b200.so (include/gnnb.h)

There is no CPU fallback: without the built library and an sm_100 GPU every operator raises.
"""
from . import _lib, ops  # noqa: F401
from .data import Data, GenotypeTable, GraphDataset  # noqa: F401

__version__ = "0.1.0"
