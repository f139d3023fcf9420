"""B200-native (sm_100a) GraphCNNDropEdge hot path -- drop-in for hoangthanh283/graph-representation-learning.

Public surface (mirrors the reference's ``gnn.models`` / ``gnn.cl_warper`` names for this path):
``GraphCNNDropEdge``, ``GraphConv``, ``NodeSelfAtten``, ``RanPACLayer``, ``make_linear_relu``, ``BaseNetwork``,
``PackedGraphBatch`` (device CSR/CSC batch), plus the collate / trainer / wrapper modules.
"""
from ._lib import GrlError, LIB_PATH  # noqa: F401
from .graph import PackedGraphBatch  # noqa: F401
from .features import SparseNodeFeatures  # noqa: F401
from .networks import (BaseNetwork, GraphCNNDropEdge, GraphConv, NodeSelfAtten, RanPACLayer,  # noqa: F401
                       RobustGCN, RPGraphCNNDropEdge, DeepRPRobustGCN, DeepRPGCN, make_linear_relu)

from .cl_warper import GNNLearningWarper  # noqa: F401,E402
from .collate import EdgeListCollate, NumpyPadding  # noqa: F401,E402

__all__ = ["GNNLearningWarper", "NumpyPadding", "EdgeListCollate", "GraphCNNDropEdge", "RobustGCN", "RPGraphCNNDropEdge", "DeepRPRobustGCN", "DeepRPGCN", "SparseNodeFeatures", "GraphConv", "NodeSelfAtten", "RanPACLayer", "make_linear_relu", "BaseNetwork",
           "PackedGraphBatch", "GrlError", "LIB_PATH"]
