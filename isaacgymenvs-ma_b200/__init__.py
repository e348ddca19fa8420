"""B200-native post-physics step kernels for IsaacGymEnvs-MA (hot path only)."""
from . import synthetic  # noqa: F401

__version__ = "0.1.0"
from . import factory  # noqa: E402,F401
from .factory import make  # noqa: E402,F401
from . import distributed  # noqa: E402,F401
from . import host_pipeline  # noqa: E402,F401
from . import rollout  # noqa: E402,F401
from . import integration  # noqa: E402,F401
from . import dr_noise  # noqa: E402,F401
