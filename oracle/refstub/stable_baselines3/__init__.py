"""TEST INFRASTRUCTURE -- import stand-in for stable-baselines3 (not installable here): user_controller/__init__.py imports
RLController, which imports PPO at module level.  Nothing on the golden generators' path instantiates it."""


class PPO:
    def __init__(self, *a, **k):
        raise NotImplementedError("stable-baselines3 is not available on this box")

    @classmethod
    def load(cls, *a, **k):
        raise NotImplementedError("stable-baselines3 is not available on this box")
