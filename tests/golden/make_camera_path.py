"""Extracts the first 1000 poses (position, camera-to-world rotation) of the reference's cameras.json (the camera path
of BASELINE configs[3]) into tests/golden/camera_path_1000.npz, float32. Data only; run in the build container:

    python tests/golden/make_camera_path.py
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
with open("/root/reference/cameras.json") as fh:
    cams = json.load(fh)[:1000]
np.savez_compressed(os.path.join(HERE, "camera_path_1000.npz"),
                    position=np.array([c["position"] for c in cams], np.float32),
                    rotation=np.array([c["rotation"] for c in cams], np.float32),
                    id=np.array([c["id"] for c in cams], np.int32))
print("wrote", len(cams), "poses")
