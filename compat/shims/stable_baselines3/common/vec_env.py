class DummyVecEnv:
    def __init__(self, *a, **k):
        raise NotImplementedError("stable-baselines3 is not installed")
