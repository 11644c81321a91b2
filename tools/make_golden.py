"""Generate tests/golden/*.npz from the reference's example data (run in the build container,
where /root/reference is mounted; the GPU box only sees the committed .npz files).

What is stored is the *input* each reference driver derives from its data files, plus the
reference's recorded final cost for (oIter, nIter, lIter) = (1, 1, 1)
(examples/test_final_cost.py:56-66); no reference source is copied.
"""
import os
import struct
import sys

import numpy as np
from PIL import Image

REF = "/root/reference/examples"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def image_warping():
    # examples/image_warping/src/main.cpp:76-81,113-127 : R channel of cat512_mask.png, 9 markers
    mask = np.array(Image.open(f"{REF}/data/cat512_mask.png"))[..., 0]
    assert set(np.unique(mask)) <= {0, 255}
    toks = open(f"{REF}/data/cat512.constraints").read().split()
    n = int(toks[0])
    cons = np.array([int(t) for t in toks[1:1 + 4 * n]], dtype=np.int32).reshape(n, 4)
    np.savez_compressed(f"{OUT}/image_warping_cat512.npz",
                        mask_is_zero=np.packbits(mask == 0), shape=np.array(mask.shape),
                        constraints=cons, golden_cost=np.float64(1774.3405))


def read_ply_vertices(path):
    with open(path, "rb") as f:
        nv, props = 0, []
        while True:
            line = f.readline().decode().strip()
            if line.startswith("element vertex"):
                nv = int(line.split()[-1]); in_v = True
            elif line.startswith("element"):
                in_v = False
            elif line.startswith("property") and in_v:
                props.append(line.split()[-1])
            elif line == "end_header":
                break
        data = np.frombuffer(f.read(nv * 4 * len(props)), dtype="<f4").reshape(nv, len(props))
    return data[:, :3].copy()


def volumetric():
    # examples/volumetric_mesh_deformation/src/CombinedSolver.h:110-121 only needs the bounding box
    v = read_ply_vertices(f"{REF}/data/head.ply")
    np.savez_compressed(f"{OUT}/volumetric_head.npz", bbox_min=v.min(0), bbox_max=v.max(0),
                        grid=np.array([5, 20, 5]), golden_cost=np.float64(189.74081))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    image_warping()
    volumetric()
    print("wrote", sorted(os.listdir(OUT)))
