"""How many rectangles of one HV level share a shape (lena 512x512, the parameters of profiles/hv_once.py)?
The search of a rectangle costs 4 * positions * pixels ~ 4 * H * W multiply-adds whatever its shape, so the question is how
much of that could be shared between rectangles of equal shape that are searched in the same launch."""
import collections
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"]
nodes, vals = fic_b200.hv_encode(img, 4, 4, 16, 16, 5, 7, 25.0)
rects, depth = fic_b200.hv_layout(nodes, 512, 512)
rects = np.asarray(rects); depth = np.asarray(depth)
w, h = rects[:, 2], rects[:, 3]
searched = ((w <= 16) & (h <= 16)) | ((w <= 4) & (h <= 4))
print("nodes", len(nodes), "leaves", int((nodes[:, 0] == 1).sum()), "searched", int(searched.sum()), "levels", int(depth.max()) + 1)
per_level = collections.Counter(depth[searched].tolist())
groups = collections.Counter(zip(depth[searched].tolist(), w[searched].tolist(), h[searched].tolist()))
sizes = np.array(list(groups.values()))
print("searched rectangles per level (top 12):", sorted(per_level.items(), key=lambda kv: -kv[1])[:12])
print("(level, shape) groups:", len(groups), "mean size", sizes.mean(), "rectangle-weighted mean size", (sizes * sizes).sum() / sizes.sum())
print("share of rectangles in groups of >= 2 / 4 / 8 / 16:", [float((sizes[sizes >= k]).sum() / sizes.sum()) for k in (2, 4, 8, 16)])
shapes = collections.Counter(zip(w[searched].tolist(), h[searched].tolist()))
print("distinct shapes over the whole tree:", len(shapes), "largest:", shapes.most_common(8))
