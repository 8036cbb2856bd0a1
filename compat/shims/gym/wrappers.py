"""The wrappers the reference applies around its environment (reference
src/rlsp/agents/simple_ddpg.py:51-98), with gym 0.26.2's single-environment behaviour."""
import time
from collections import deque

import numpy as np

from .core import Wrapper


class AutoResetWrapper(Wrapper):
    def step(self, action):
        obs, reward, terminated, truncated, info = self.env.step(action)
        if terminated or truncated:
            new_obs, new_info = self.env.reset()
            new_info["final_observation"] = obs
            new_info["final_info"] = info
            obs, info = new_obs, new_info
        return obs, reward, terminated, truncated, info


class RecordEpisodeStatistics(Wrapper):
    def __init__(self, env, deque_size=100):
        super().__init__(env)
        self.num_envs = 1
        self.is_vector_env = False
        self.t0 = time.perf_counter()
        self.episode_count = 0
        self.episode_returns = None
        self.episode_lengths = None
        self.return_queue = deque(maxlen=deque_size)
        self.length_queue = deque(maxlen=deque_size)

    def reset(self, **kwargs):
        obs = self.env.reset(**kwargs)
        self.episode_returns = np.zeros(self.num_envs, dtype=np.float32)
        self.episode_lengths = np.zeros(self.num_envs, dtype=np.int32)
        return obs

    def step(self, action):
        obs, reward, terminated, truncated, info = self.env.step(action)
        assert isinstance(info, dict)
        self.episode_returns += reward
        self.episode_lengths += 1
        if terminated or truncated:
            episode_return, episode_length = self.episode_returns[0], self.episode_lengths[0]
            info = {**info, "episode": {"r": episode_return, "l": episode_length,
                                         "t": round(time.perf_counter() - self.t0, 6)}}
            self.return_queue.append(episode_return)
            self.length_queue.append(episode_length)
            self.episode_count += 1
            self.episode_returns[0] = 0
            self.episode_lengths[0] = 0
        return obs, reward, terminated, truncated, info


class RecordVideo(Wrapper):
    def __init__(self, env, *a, **k):
        raise NotImplementedError("RecordVideo needs a rendering stack; the reference never enables it")
