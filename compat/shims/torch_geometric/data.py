from gsc_b200.graph import GraphBatch as Batch  # noqa: F401
from gsc_b200.graph import GraphData as Data  # noqa: F401
