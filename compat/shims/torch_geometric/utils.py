"""``from_networkx`` as PyG 2.3.0 defines it, for the arguments the reference passes
(``group_node_attrs=[...]``, ``group_edge_attrs=None``; reference simulator_wrapper.py:297-299): nodes are
relabelled 0..n-1 in ``G.nodes`` order, an undirected graph lists both directions of every link, ordered by
source node and then by adjacency order; the grouped node attributes become the columns of ``x``."""
import networkx as nx
import torch

from gsc_b200.graph import GraphData


def from_networkx(G, group_node_attrs=None, group_edge_attrs=None):
    if group_edge_attrs:
        raise NotImplementedError("group_edge_attrs is not used by the reference")
    G = nx.convert_node_labels_to_integers(G)
    G = G.to_directed() if not nx.is_directed(G) else G
    edges = list(G.edges(keys=False)) if isinstance(G, (nx.MultiGraph, nx.MultiDiGraph)) else list(G.edges)
    edge_index = torch.tensor(edges, dtype=torch.long).t().contiguous().view(2, -1)
    attrs = list(group_node_attrs or [])
    cols = []
    for key in attrs:
        col = torch.tensor([feat[key] for _, feat in G.nodes(data=True)])
        cols.append(col.view(-1, 1) if col.dim() <= 1 else col)
    if cols:
        x = torch.cat(cols, dim=-1)
    else:
        x = torch.zeros(G.number_of_nodes(), 0)
    return GraphData(x, edge_index)
