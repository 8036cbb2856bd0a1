"""gsc_b200 -- B200-native hot path of GSC (GNN embedder + DDPG actor/critic), see DESIGN.md."""
from .graph import GraphBatch, GraphData  # noqa: F401

__all__ = ["GraphBatch", "GraphData"]
