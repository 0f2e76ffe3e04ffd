"""Extracts planner trajectories logged by the reference PROJECT (Blender/scenes with animation/*.json) into
tests/golden/reference_logged_paths.json.gz.

Run in the build container (needs /root/reference):  python tests/golden/make_logged_paths.py

What these are, and are not: the files carry the keys `path_to_strrt_config_json`, `execution_time_ms`,
`number_of_state_validations`, `path: [[t, [q...]], ...]` -- the writer of the legacy Python prototype
(STRRT/STRRT_eval/evaluate_blender_scene.py:414-422, collision checker: pybullet), NOT the keys of the shipped C++
writer (ResultsWriter.cpp:69-99: `path_to_config_json`, `full_execution_time_ns`, `final_planner_data`).  They are
therefore outputs of an OLDER pipeline with a DIFFERENT checker, on the same scenes and robot model.  They do not pin
the shipped checker (CollisionManager.cpp:330-380); they are a centimetre-level sanity check of FK, hulls, box extents
and pose tables: a plausible collision-free trajectory that passes 2-31 mm from the obstacles must stay collision-free
in our oracle and on the GPU.  Only data is extracted (time stamps and joint vectors); no reference source is copied."""
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
