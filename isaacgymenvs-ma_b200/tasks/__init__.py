"""Host-side mirrors of the reference task classes for the post-physics hot path.

Same attribute names, buffer shapes / dtypes and ``step()`` contract as
``isaacgymenvs/tasks/base/vec_task.py`` and the task modules; ``post_physics_step`` is one call
into the C ABI (libige_b200.so).  ``task_map`` mirrors ``isaacgym_task_map``
(isaacgymenvs/tasks/__init__.py:67-100) for the tasks on the path.
"""
from .base import SyntheticSim, VecTaskB200  # noqa: F401
from .cartpole import Cartpole  # noqa: F401
from .locomotion import Ant, Humanoid  # noqa: F401
from .shadow_hand import AllegroHand, ShadowHand  # noqa: F401
from .anymal import Anymal  # noqa: F401
from .small import BallBalance, Ingenuity, Quadcopter  # noqa: F401
from .franka_sa import FrankaCabinet, FrankaCubeStack, FrankaCubeStack2, FrankaReach  # noqa: F401
from .franka_ma import FrankaCollectMA, FrankaCombineMA, FrankaPPMA, FrankaReachMA  # noqa: F401

task_map = {"Cartpole": Cartpole, "Ant": Ant, "Humanoid": Humanoid, "ShadowHand": ShadowHand,
            "FrankaReachMA": FrankaReachMA, "FrankaCollectMA": FrankaCollectMA, "FrankaCombineMA": FrankaCombineMA,
            "FrankaPPMA": FrankaPPMA, "Anymal": Anymal, "BallBalance": BallBalance, "Quadcopter": Quadcopter,
            "Ingenuity": Ingenuity, "FrankaReach": FrankaReach, "FrankaCubeStack2": FrankaCubeStack2,
            "FrankaCubeStack": FrankaCubeStack, "FrankaCabinet": FrankaCabinet, "AllegroHand": AllegroHand}
