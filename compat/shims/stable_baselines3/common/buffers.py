class _Unavailable:
    def __init__(self, *a, **k):
        raise NotImplementedError("stable-baselines3 is not installed: only graph_mode=True (GraphReplayBuffer) is supported")


class ReplayBuffer(_Unavailable):
    pass


class DictReplayBuffer(_Unavailable):
    pass
