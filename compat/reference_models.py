"""What ``src/rlsp/agents/models.py`` becomes when the reference is run on gsc_b200.

INTEGRATION.md section 3 replaces that file by ``from gsc_b200.models import *``.  The reference pins its
device to ``"cpu"`` (reference src/rlsp/agents/simple_ddpg.py:106-108) and keeps every observation and loss
on the host, while gsc_b200 has no CPU path.  The two subclasses below bridge that without touching the
reference: ``.to("cpu")`` leaves the parameters on the GPU, ``forward`` moves the observation (and the
action) to the GPU and hands the result back on the caller's device.  Both copies are differentiable, so
``loss.backward()`` and the reference's two ``optim.Adam`` instances work unchanged."""
import torch

from gsc_b200 import models as _m
from gsc_b200.graph import GraphBatch
from gsc_b200.models import GNNEmbedder  # noqa: F401  (same public names as the reference module)

_DEVICE = torch.device("cuda", 0)


class _GpuResident:
    def to(self, *args, **kwargs):
        device, dtype, non_blocking, _ = torch._C._nn._parse_to(*args, **kwargs)
        if device is not None and device.type == "cpu":
            device = _DEVICE                        # the hot path has no CPU implementation
        return super().to(device if device is not None else _DEVICE, dtype, non_blocking)

    def _to_gpu(self, obs):
        gb = GraphBatch.wrap(obs)
        return gb, gb.x.device


class Actor(_GpuResident, _m.Actor):
    def __init__(self, agent_helper):
        super().__init__(agent_helper)
        self.to(_DEVICE)

    def forward(self, x):
        gb, dev = self._to_gpu(x)
        return super().forward(gb.to(self.flat.device)).to(dev)


class QNetwork(_GpuResident, _m.QNetwork):
    def __init__(self, agent_helper):
        super().__init__(agent_helper)
        self.to(_DEVICE)

    def forward(self, x, a):
        gb, dev = self._to_gpu(x)
        return super().forward(gb.to(self.flat.device), a.to(self.flat.device)).to(dev)
