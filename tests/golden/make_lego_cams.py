"""Extracts a few camera-to-world matrices of the NeRF-synthetic lego set the reference ships
(/root/reference/nerf_synthetic/lego/transforms_train.json) into tests/golden/lego_cams.json so the
config-1 shaped tests can run where /root/reference does not exist.  Run in the build container:
    python tests/golden/make_lego_cams.py
"""
import json
import os

SRC = "/root/reference/nerf_synthetic/lego/transforms_train.json"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lego_cams.json")

if __name__ == "__main__":
    with open(SRC) as f:
        t = json.load(f)
    frames = t["frames"]
    keep = [0, 7, 23, 42, 64, 99]
    out = {"camera_angle_x": t["camera_angle_x"], "source": SRC, "frame_indices": keep,
           "transform_matrix": [frames[i]["transform_matrix"] for i in keep]}
    with open(OUT, "w") as f:
        json.dump(out, f)
    print("wrote", OUT, len(keep), "cameras, camera_angle_x =", t["camera_angle_x"])
