"""Graph replay buffer with the reference's surface (reference src/rlsp/agents/buffer.py:8-89).

``add`` stores one transition, ``sample`` draws indices with the same numpy global-RNG calls as the
reference (:57-62) and returns a ``BufferSample`` whose observation fields are ``GraphBatch`` objects
in ``Batch.from_data_list`` layout, already on ``device``; their CSR is packed on the GPU the first
time a model touches them.  Host-side storage stays Python, as in the reference.
"""
from __future__ import annotations

from typing import NamedTuple

import numpy as np
import torch

from .graph import GraphBatch, GraphData


class BufferSample(NamedTuple):
    observations: GraphBatch
    actions: torch.Tensor
    next_observations: GraphBatch
    dones: torch.Tensor
    rewards: torch.Tensor


class GraphReplayBuffer:
    def __init__(self, buffer_size: int, action_space, device="cpu") -> None:
        self.buffer_size = int(buffer_size)
        self.device = device
        self.action_space = int(np.prod(action_space.shape))
        self.dones = np.ones(self.buffer_size, np.float32)
        self.rewards = np.zeros(self.buffer_size, dtype=np.float32)
        self.actions = torch.zeros((self.buffer_size, self.action_space), dtype=torch.float32)
        self.observations = [None] * self.buffer_size
        self.next_observations = [None] * self.buffer_size
        self.pos = 0
        self.full = False

    @staticmethod
    def _as_data(obs) -> GraphData:
        return obs.clone() if isinstance(obs, GraphData) else GraphData(obs.x.clone(), obs.edge_index.clone(),
                                                                        obs.mask.clone())

    def add(self, obs, next_obs, action: np.ndarray, reward, done, infos=None) -> None:
        self.dones[self.pos] = done
        self.rewards[self.pos] = reward
        self.actions[self.pos] = torch.from_numpy(np.asarray(action)).to(torch.float32)
        self.observations[self.pos] = self._as_data(obs)
        self.next_observations[self.pos] = self._as_data(next_obs)
        self.pos += 1
        if self.pos == self.buffer_size:
            self.pos = 0
            self.full = True

    def sample(self, batch_size: int) -> BufferSample:
        if self.full:
            batch_inds = (np.random.randint(1, self.buffer_size, size=batch_size) + self.pos) % self.buffer_size
        else:
            batch_inds = np.random.randint(0, self.pos, size=batch_size)
        cur = GraphBatch.from_data_list([self.observations[i] for i in batch_inds]).to(self.device)
        nxt = GraphBatch.from_data_list([self.next_observations[i] for i in batch_inds]).to(self.device)
        cur.mask = cur.mask.view(-1, self.action_space)
        nxt.mask = nxt.mask.view(-1, self.action_space)
        return BufferSample(
            observations=cur,
            actions=self.actions[batch_inds].to(self.device),
            next_observations=nxt,
            dones=torch.from_numpy(self.dones[batch_inds]).to(self.device),
            rewards=torch.from_numpy(self.rewards[batch_inds]).to(self.device),
        )
