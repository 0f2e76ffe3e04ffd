"""Regenerates tests/golden/* from the reference's shipped scenes and robot models.

Run in the build container (needs /root/reference; the GPU box does not have it):
    python tests/golden/make_golden.py

Outputs
  *.pack.gz            scene packs (include/sirrt_pack.h) of the reference's scenes / robot models
  golden_vectors.json  per scene: configurations and the ORACLE's safe intervals, per-frame checks
                       and FK link origins for them.

What pins what: the packs are produced from the reference's own JSON/URDF/STL inputs by
tools/scene_pack.py; the vectors are outputs of oracle/sirrt_oracle.c (the reference binary cannot
be built in this image: coal, Eigen, nanoflann, urdfdom, orocos-kdl are absent), so they are
regression vectors for the oracle and parity targets for the CUDA path, not reference outputs.
The only reference-derived known answers are the check_scene invariants (start free at frame 0,
goal free at the last frame; STRRT_Planner/src/util_executables/check_scene.cpp:51-103) and the FK
sanity values of SURVEY.md §8c, both asserted in tests/test_oracle.py.
"""
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

REF = "/root/reference"
SCENES = {
    "mass_test": "STRRT/mass_test.json",
    "flying_boxes": "Blender/scenes/scene_task_flying_boxes.json",
    "multiple_robots": "STRRT/scene_task_multiple_robots.json",
    "scene_task": "STRRT/scene_task.json",
}


def main():
    sp, orc = g.load_tools(), g.load_oracle()
    orc.build()
    vectors = {}
    rng = np.random.default_rng(20260101)
    for name, rel in SCENES.items():
        scene = sp.load_scene_json(os.path.join(REF, rel), REF)
        pack = sp.write_pack(scene)
        with gzip.GzipFile(os.path.join(HERE, name + ".pack.gz"), "wb", compresslevel=9, mtime=0) as f:
            f.write(pack)
        o = orc.Oracle(pack, orc.REFERENCE_EXACT)
        lo, hi = scene.robot.limits[:, 0].astype(float), scene.robot.limits[:, 1].astype(float)
        qs = [scene.start_configuration, scene.end_configuration] + [lo + rng.random(o.dof) * (hi - lo) for _ in range(6)]
        ent = {"source": rel, "n_frames": o.n_frames, "configs": [], "intervals": [], "frame_checks": [], "fk_origins": []}
        for q in qs:
            q = np.asarray(q, np.float64)
            ent["configs"].append([float(x) for x in q])
            ent["intervals"].append(o.safe_intervals(q))
            fr = [0, o.n_frames // 3, o.n_frames - 1]
            ent["frame_checks"].append([[f, int(o.check_frame(q, f))] for f in fr])
            ent["fk_origins"].append([[float(x) for x in t] for t in o.fk(q)[1]])
        vectors[name] = ent
        print(name, len(pack), "bytes;", [len(iv) for iv in ent["intervals"]], "intervals")
    # the capsule model of the xArm6 (multi-agent generators use it): robot-only pack
    cyl = sp.parse_urdf(os.path.join(REF, "Blender/robots/xarm6/xarm6cyl.urdf"), [f"joint{i}" for i in range(1, 7)], [0, 0, 0], [0, 0, 0, 1])
    z = np.zeros(6)
    pack = sp.write_pack(sp.Scene(1, 30, 0.0, z, z, np.full(6, np.pi), cyl))
    with gzip.GzipFile(os.path.join(HERE, "xarm6cyl_robot.pack.gz"), "wb", compresslevel=9, mtime=0) as f:
        f.write(pack)
    with open(os.path.join(HERE, "golden_vectors.json"), "w") as f:
        json.dump(vectors, f, separators=(",", ":"))
    print("done")


if __name__ == "__main__":
    main()
