"""Derive a cross-check fixture for the marching-cubes restatement from the only MC tables in the reference tree
(build container only):  third_party/colmap/lib/PoissonRecon/MarchingCubes.cpp  (classic 256-case `triangles` table,
lines 528-785; index convention `GetIndex`, lines 789-801; edge convention `Cube::EdgeCorners`, lines 194-212).

    python -m oracle.make_mc_classic_golden      ->  tests/golden/mc_classic_cases.npz

TEST INFRASTRUCTURE.  HIVE does not use these tables (it calls skimage's Lewiner implementation, which is not on the
box: parity unpinned), but for every sign configuration WITHOUT an ambiguous face the Lewiner tiling and the classic one
describe the same surface, so triangle counts and the split into connected patches must agree.  Stored per
configuration (numbered by our own convention: bit i <=> corner (i&1, (i>>1)&1, (i>>2)&1) is *below* the level):
``ntri`` and ``patches`` (sorted triangle counts of the connected patches, zero padded).  No table text is copied.
"""
from __future__ import annotations

import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

# bit b of PoissonRecon's mcIndex <-> corner position (x, y, z)   (MarchingCubes.cpp:789-801)
PR_BIT_POS = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]


def parse_triangles(path):
    text = open(path).read()
    start = text.index("MarchingCubes::triangles[")
    body = text[text.index("{", start) + 1:]
    rows = re.findall(r"\{([^{}]*)\}", body)[:256]
    table = [[int(v) for v in r.replace("\n", " ").split(",") if v.strip()] for r in rows]
    assert len(table) == 256 and all(len(r) == 16 for r in table)
    return table


def patches_of(tris):
    """Connected patches (triangles sharing a cube edge index) -> sorted sizes."""
    parent = list(range(len(tris)))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a
    for i in range(len(tris)):
        for j in range(i):
            if len(set(tris[i]) & set(tris[j])) >= 1:  # patches of one cube never share an edge-vertex
                parent[find(i)] = find(j)
    sizes = {}
    for i in range(len(tris)):
        sizes[find(i)] = sizes.get(find(i), 0) + 1
    return sorted(sizes.values(), reverse=True)


def main():
    assert ref_shim.available()
    src = os.path.join(ref_shim.REFERENCE_ROOT, "third_party", "colmap", "lib", "PoissonRecon", "MarchingCubes.cpp")
    table = parse_triangles(src)
    ntri = np.zeros(256, np.int32)
    patches = np.zeros((256, 5), np.int32)
    for ours in range(256):
        below = {(i & 1, (i >> 1) & 1, (i >> 2) & 1) for i in range(8) if (ours >> i) & 1}
        pr = sum(1 << b for b, pos in enumerate(PR_BIT_POS) if pos in below)
        row = [v for v in table[pr] if v >= 0]
        tris = [tuple(row[k:k + 3]) for k in range(0, len(row), 3)]
        ntri[ours] = len(tris)
        p = patches_of(tris)
        patches[ours, :len(p)] = p
    out = os.path.join(ROOT, "tests", "golden", "mc_classic_cases.npz")
    np.savez_compressed(out, ntri=ntri, patches=patches)
    print(out, os.path.getsize(out), "bytes; triangle-count histogram", np.bincount(ntri))


if __name__ == "__main__":
    main()
