"""gnn-cca_b200: the GNN-CCA message-passing hot path on B200 (sm_100a).

Import with ``importlib.import_module("gnn-cca_b200")`` or through the alias module ``gnncca`` at the
repository root.  Public surface mirrors the reference (vpulab/GNN-CCA ``models/mpn.py``, ``models/mlp.py``)
plus the graph construction that the reference writes inline in its loops."""
from .mlp import MLP
from .mpn import MOTMPNet, MetaLayer, EdgeModel, NodeModel, MLPGraphIndependent
from .plan import GraphPlan
from .graph import Batch, build_graph_batch, column_normalize
from . import _lib
from .engine import forward_host, StageTimer, HostPipeline
from . import multigpu
from .loss import compute_loss_acc, loss_stats, LossStats
from . import postprocess
from .graphs import GraphedTrainStep
from . import detections
from .postprocess import decide_and_cluster, ground_truth_clusters, frame_clusters

__all__ = ["MLP", "MOTMPNet", "MetaLayer", "EdgeModel", "NodeModel", "MLPGraphIndependent", "GraphPlan", "Batch", "build_graph_batch", "column_normalize", "forward_host", "StageTimer", "HostPipeline", "compute_loss_acc", "loss_stats", "LossStats", "decide_and_cluster", "ground_truth_clusters", "frame_clusters", "GraphedTrainStep"]
