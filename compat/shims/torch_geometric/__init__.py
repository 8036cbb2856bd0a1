"""Stand-in for the parts of torch-geometric 2.3.0 that the reference's HOST code touches: the ``Data`` /
``Batch`` containers (reference src/rlsp/agents/buffer.py:4,65-81, simple_ddpg.py:19,36-38) and
``from_networkx`` (reference src/rlsp/envs/simulator_wrapper.py:12,297).  The containers are gsc_b200's own
(`gsc_b200.graph`), which were written to this surface.  ``torch_geometric.nn`` (GATv2Conv, MLP,
global_mean_pool) is deliberately absent: the reference's models module is what gsc_b200 replaces."""
from . import data, utils  # noqa: F401

__version__ = "2.3.0-compat"
