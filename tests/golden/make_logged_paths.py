"""Extracts the planner trajectories the REFERENCE itself logged (Blender/scenes with animation/*.json:
outputs of the reference's ST-RRT* and DRGBT executables on its shipped scenes) into
tests/golden/reference_logged_paths.json.gz.

Run in the build container (needs /root/reference):  python tests/golden/make_logged_paths.py

Why these pin the oracle (one-sided, at the a4 boundary): every state of an ST-RRT* solution was accepted by
the reference's own checker -- SpaceTimeMotionValidator::checkMotion (SpaceTimeMotionValidator.cpp:22-74)
interpolates each tree edge into max(int(|dq|/0.1), int(2*dt*fps+0.5)) + 2 states and calls
StateValidityCheckerFunctor::isValid -> CollisionManager::check_collision(q, float time)
(StateValidityCheckerFunctor.cpp:29-47, CollisionManager.cpp:330-336) on each; the DRGBT logs are executed
trajectories sampled every 10 ms.  The oracle must therefore report every one of these (q, frame) cells FREE.
Only data is extracted (time stamps and joint vectors); no reference source is copied."""
import gzip
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/Blender/scenes with animation"
LOGS = {
    # key: (scene pack under tests/golden, log file, planner)
    "strrt_mass_test": ("mass_test", "mass_test_strrt_logs.json"),
    "strrt_flying_boxes": ("flying_boxes", "strrt_logs_flying_boxes.json"),
    "strrt_multiple_robots": ("multiple_robots", "strrt_logs_multiple_robot_on_table.json"),
    "drgbt_mass_test": ("mass_test", "mass test 1 50 boxes drgbt.json"),
    "rgbtstar_mass_test": ("mass_test", "mass test 1 50 boxes drgbt RGBTStar.json"),
    "drgbt_flying_boxes": ("flying_boxes", "drgbt_logs_flying_boxes.json"),
    "drgbt_multiple_robots": ("multiple_robots", "drgbt_logs_multiple_robot_on_table.json"),
}


def main():
    out = {}
    for key, (scene, fn) in LOGS.items():
        d = json.load(open(os.path.join(SRC, fn)))
        out[key] = {"scene": scene, "source": "Blender/scenes with animation/" + fn, "planner": d["planner_type"],
                    "scene_json": d["path_to_scene_json"], "t": [p[0] for p in d["path"]], "q": [p[1] for p in d["path"]]}
        print(key, d["planner_type"], len(d["path"]), "states")
    with gzip.GzipFile(os.path.join(HERE, "reference_logged_paths.json.gz"), "wb", compresslevel=9, mtime=0) as f:
        f.write(json.dumps(out, separators=(",", ":")).encode())


if __name__ == "__main__":
    main()
