#!/usr/bin/env python
"""Stage the reference's own Python sources for the hot path under the git-ignored ``oracle/_ref/``.

The reference (Xhadow0823/IsaacGymEnvs-MA) is pure Python + torch; its post-physics path needs no build.
``/root/reference`` only exists in the build container, so ``__graft_entry__.build()`` runs this recipe there
and the staged tree travels to the GPU box with the snapshot (``oracle/_ref/`` is in .gitignore, NOT in
.gpurunignore) -- exactly like a compiled ``oracle/_ref/*.so`` would for a C reference.  Nothing is copied
into git history and the files are byte-for-byte the reference's (unmodified: ``tests/ref_shim.py`` stubs
``isaacgym`` / ``gym`` around them at import time).

Consumers (checker / baseline only, never the product path):
  * ``bench.py --impl reference`` and ``bench.py``'s ``cpu_baseline`` leg: the reference's own
    ``Task.post_physics_step`` on CPU torch (``cpu_baseline.kind = "reference"``),
  * tests marked ``needs_reference`` (class-level parity against the unmodified classes).
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("IGE_REFERENCE_SRC", "/root/reference")
DST = os.path.join(ROOT, "oracle", "_ref")

# SURVEY.md section 8(a) files + what they import + the 8(f) callers
FILES = [
    "isaacgymenvs/tasks/base/__init__.py",
    "isaacgymenvs/tasks/base/vec_task.py",
    "isaacgymenvs/utils/torch_jit_utils.py",
    "isaacgymenvs/utils/dr_utils.py",
    "isaacgymenvs/utils/utils.py",
    "isaacgymenvs/learning/A2CAgent_MA.py",
] + [f"isaacgymenvs/tasks/{t}.py" for t in (
    "cartpole", "ant", "humanoid", "shadow_hand", "anymal",
    "franka_reach_MA", "franka_collect_MA", "franka_combine_MA", "franka_PPMA",
    "franka_reach", "franka_cube_stack2", "franka_cube_stack", "franka_cabinet",
    "allegro_hand", "ball_balance", "quadcopter", "ingenuity")]


def stage(verbose=False):
    """Copy FILES from SRC to DST.  Returns the number of files staged (0 when SRC is absent)."""
    if not os.path.isdir(os.path.join(SRC, "isaacgymenvs", "tasks")):
        return 0
    n = 0
    for rel in FILES:
        s, d = os.path.join(SRC, rel), os.path.join(DST, rel)
        if not os.path.exists(s):
            continue
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        n += 1
        if verbose:
            print("staged", rel)
    with open(os.path.join(DST, "README"), "w") as f:
        f.write("Unmodified files of the reference repository, staged by tools/vendor_reference.py.\n"
                "Git-ignored; test / baseline infrastructure only.\n")
    return n


if __name__ == "__main__":
    k = stage(verbose="-v" in sys.argv)
    print(f"{k} reference files staged under {DST}" if k else f"{SRC} not present: nothing staged")
