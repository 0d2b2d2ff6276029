"""Copies the one processed asset the reference ships (imported_assets/Default/box.glb, 6253 bytes; its .meta names
RayTracer3::assets::MeshLoader) into tests/golden/ref_box_mesh.npz so that tests/test_assets.py can pin the bincode element codecs
against bytes written by the reference's own MeshSaver without reading /root/reference at test time."""
import hashlib
import os

import numpy as np

SRC = "/root/reference/imported_assets/Default/box.glb"
HERE = os.path.dirname(os.path.abspath(__file__))

if __name__ == "__main__":
    raw = open(SRC, "rb").read()
    np.savez_compressed(os.path.join(HERE, "ref_box_mesh.npz"), raw=np.frombuffer(raw, np.uint8), sha256=np.array(hashlib.sha256(raw).hexdigest()))
    print(len(raw), hashlib.sha256(raw).hexdigest())
