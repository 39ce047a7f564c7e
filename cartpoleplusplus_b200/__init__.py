"""cartpoleplusplus_b200 — B200-native batched simulator for the cartpole++ env step."""
from .opts import add_opts, default_opts  # noqa: F401
from .config import PhysicsParams, make_config  # noqa: F401
