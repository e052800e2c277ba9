#!/usr/bin/env python
"""Generate the marching-cubes tables used by the CUDA kernels and by the C oracle.

    python tools/gen_mc_tables.py          # rewrites hive_b200/csrc/hv_mc_tables.h and oracle/mc_tables.h

Why generated: the reference extracts its mesh with ``skimage.measure.marching_cubes`` (Lewiner et al. 2003),
scikit-image==0.19.1, which is a third-party dependency absent from /root/reference and from this image (SURVEY.md
§8(c), App. C) -- its look-up tables cannot be consulted, and they must not be typed from memory.  What *is* published
is the construction: the surface inside a cube is bounded by the iso-contours on the six cube faces; a face whose
corners alternate in sign is ambiguous and is resolved with the asymptotic decider (the "face test" A*C - B*D of
Chernyaev / Lewiner); the contour segments chain into closed loops and each loop bounds a disk.  This script derives
the tables from exactly that construction:

  * corner / edge numbering and the case index follow the skimage Cython (App. C): bit i set iff v_i > level;
  * for each of the 256 sign configurations and each resolution of its ambiguous faces ("variant"), the face segments
    are oriented, chained into loops, and every loop is triangulated without chords that lie in a cube face (such a
    chord could coincide with one of the neighbouring cube: a non-manifold edge); the 8-, 9- and 12-gons that admit no
    such triangulation are fanned around an extra centre vertex, as Lewiner's tilings 7.3, 10.2, 12.2, 13.x do;
  * orientation: right-hand normals point towards increasing values in the (axis0, axis1, axis2) output frame, which
    is what skimage returns for gradient_direction='descent' (its default; App. C).

Scope note (DESIGN.md "Marching cubes"): the interior ("tunnel") sub-cases of Lewiner's cases 4, 6, 7, 10, 12, 13 are
tiled as disks, i.e. the result is topologically consistent across cubes (watertight) but an isolated cube may hold two
disks where the trilinear interpolant has a tube.
"""
from __future__ import annotations

import itertools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# Corner i -> offsets along (axis0, axis1, axis2).  skimage's Cython frame is x=axis2, y=axis1, z=axis0 with
# v0(0,0,0) v1(1,0,0) v2(1,1,0) v3(0,1,0) v4(0,0,1) v5(1,0,1) v6(1,1,1) v7(0,1,1) in (x,y,z).
CORNERS = np.array([(0, 0, 0), (0, 0, 1), (0, 1, 1), (0, 1, 0), (1, 0, 0), (1, 0, 1), (1, 1, 1), (1, 1, 0)])
# Edge e -> its two corners (Lewiner / skimage numbering: 0-3 bottom ring, 4-7 top ring, 8-11 verticals).
EDGES = [(0, 1), (1, 2), (2, 3), (3, 0), (4, 5), (5, 6), (6, 7), (7, 4), (0, 4), (1, 5), (2, 6), (3, 7)]


def edge_owner(e):
    """(offset of the owning voxel, axis) of edge e: a lattice edge belongs to its lower corner."""
    a, b = CORNERS[EDGES[e][0]], CORNERS[EDGES[e][1]]
    lo = np.minimum(a, b)
    axis = int(np.nonzero(a != b)[0][0])
    return tuple(int(v) for v in lo), axis


def _faces():
    faces = []
    for axis in range(3):
        for side in (0, 1):
            idx = [i for i in range(8) if CORNERS[i][axis] == side]
            normal = np.zeros(3)
            normal[axis] = 1.0 if side else -1.0
            centre = CORNERS[idx].mean(axis=0)
            # cyclic order: sort by angle in the face plane
            u_axis, v_axis = [k for k in range(3) if k != axis]
            ang = [np.arctan2(CORNERS[i][v_axis] - centre[v_axis], CORNERS[i][u_axis] - centre[u_axis]) for i in idx]
            order = [i for _, i in sorted(zip(ang, idx))]
            faces.append((order, normal))
    return faces


FACES = _faces()  # 6 x (4 corners in cyclic order, outward normal)
EDGE_OF = {frozenset(c): e for e, c in enumerate(EDGES)}
EDGE_MID = np.array([(CORNERS[a] + CORNERS[b]) / 2.0 for a, b in EDGES])


def ambiguous_faces(config):
    """Indices of the faces whose corners alternate in sign."""
    out = []
    for f, (cyc, _) in enumerate(FACES):
        s = [(config >> c) & 1 for c in cyc]
        if s[0] == s[2] and s[1] == s[3] and s[0] != s[1]:
            out.append(f)
    return out


def face_segments(config, f, connected):
    """Oriented contour segments (edge_from, edge_to) of face f.  `connected`: the positive diagonal pair of an
    ambiguous face is joined through the face centre (asymptotic decider positive)."""
    cyc, normal = FACES[f]
    s = [(config >> c) & 1 for c in cyc]
    n_pos = sum(s)
    if n_pos in (0, 4):
        return []
    groups = []  # (corner positions cut off by the segment, sign of that group)
    if n_pos == 1 or n_pos == 3:
        odd = s.index(1 if n_pos == 1 else 0)
        groups.append(([odd], s[odd]))
    elif s[0] == s[2]:  # ambiguous
        want = 0 if connected else 1  # connected positives: the segments cut off the two negative corners
        for k in range(4):
            if s[k] == want:
                groups.append(([k], want))
    else:  # two adjacent positives
        k = [i for i in range(4) if s[i] == 1 and s[(i + 1) % 4] == 1][0]
        groups.append(([k, (k + 1) % 4], 1))
    segs = []
    for pos, sign in groups:
        first, last = pos[0], pos[-1]
        e_a = EDGE_OF[frozenset((cyc[(first - 1) % 4], cyc[first]))]
        e_b = EDGE_OF[frozenset((cyc[last], cyc[(last + 1) % 4]))]
        p, q = EDGE_MID[e_a], EDGE_MID[e_b]
        centroid = CORNERS[[cyc[k] for k in pos]].mean(axis=0)
        side = (centroid - (p + q) / 2.0) * (1.0 if sign else -1.0)  # towards the positive side of the segment
        if np.dot(np.cross(normal, q - p), side) > 0:
            segs.append((e_a, e_b))
        else:
            segs.append((e_b, e_a))
    return segs


CENTRE = 12  # pseudo-edge: the extra vertex inside the cube (Lewiner's "c" vertex)
_FACES_OF_EDGE = {e: set() for e in range(12)}
for _f, (_cyc, _) in enumerate(FACES):
    for _k in range(4):
        _FACES_OF_EDGE[EDGE_OF[frozenset((_cyc[_k], _cyc[(_k + 1) % 4]))]].add(_f)


def _coplanar(e_a, e_b):
    """Both edge vertices lie on a common cube face: a chord between them would lie IN that face and could coincide
    with a chord of the neighbouring cube (a non-manifold edge), so only contour segments may join such a pair."""
    return bool(_FACES_OF_EDGE[e_a] & _FACES_OF_EDGE[e_b])


def cube_loops(config, face_bits=0):
    """Closed, oriented contour loops (lists of edge indices) of one cube."""
    amb = ambiguous_faces(config)
    nxt = {}
    for f in range(6):
        connected = bool((face_bits >> amb.index(f)) & 1) if f in amb else False
        for a, b in face_segments(config, f, connected):
            assert a not in nxt, "an edge vertex has two outgoing segments"
            nxt[a] = b
    loops, seen = [], set()
    for start in sorted(nxt):
        if start in seen:
            continue
        loop, e = [], start
        while e not in seen:
            seen.add(e)
            loop.append(e)
            e = nxt[e]
        assert e == start and len(loop) >= 3
        loops.append(loop)
    crossing = [e for e, (a, b) in enumerate(EDGES) if ((config >> a) & 1) != ((config >> b) & 1)]
    assert sorted(seen) == crossing
    return loops


def _triangulate(loop):
    """First (in a fixed enumeration order) triangulation of the loop without in-face chords; None if there is none."""
    memo = {}

    def rec(i, j):
        if j - i < 2:
            return ()
        if (i, j) not in memo:
            found = None
            for k in range(i + 1, j):
                if (k != i + 1 and _coplanar(loop[i], loop[k])) or (j != k + 1 and _coplanar(loop[k], loop[j])):
                    continue
                left, right = rec(i, k), rec(k, j)
                if left is None or right is None:
                    continue
                found = left + ((loop[i], loop[k], loop[j]),) + right
                break
            memo[(i, j)] = found
        return memo[(i, j)]

    return rec(0, len(loop) - 1)


def tile_cube(config, face_bits=0):
    """Triangles (edge-index triples, 12 = centre vertex) of one cube.  Bit k of `face_bits` resolves the k-th
    ambiguous face.  Every loop is triangulated without in-face chords; where no such triangulation exists
    (the 8-, 9- and 12-gons of Lewiner's cases 10/12, 7.3 and 13) the loop is fanned around the centre vertex."""
    tris, used_centre = [], False
    for loop in cube_loops(config, face_bits):
        t = _triangulate(loop)
        if t is None:
            assert not used_centre, "two loops of one cube need the centre vertex"
            used_centre = True
            t = tuple((CENTRE, loop[i], loop[(i + 1) % len(loop)]) for i in range(len(loop)))
        tris.extend(t)
    return tris


def build_tables():
    amb_count, amb_faces, var_base = [], [], []
    var_ntri, var_off, var_mult, tris = [], [], [], []
    for config in range(256):
        amb = ambiguous_faces(config)
        amb_count.append(len(amb))
        amb_faces.append(sum(f << (3 * k) for k, f in enumerate(amb)))
        var_base.append(len(var_ntri))
        for bits in range(1 << len(amb)):
            t = tile_cube(config, bits)
            var_ntri.append(len(t))
            var_off.append(len(tris))
            mult = [0] * 13
            for tri in t:
                for e in tri:
                    mult[e] += 1
            assert max(mult) < 16
            var_mult.append(sum(m << (4 * e) for e, m in enumerate(mult)))
            tris.extend(t)
    return dict(amb_count=amb_count, amb_faces=amb_faces, var_base=var_base, var_ntri=var_ntri, var_off=var_off,
                var_mult=var_mult, tris=tris)


def _c_array(ctype, name, values, per_line=16, fmt="%d"):
    lines = [f"HV_MC_TABLE {ctype} {name}[{len(values)}] = {{"]
    for i in range(0, len(values), per_line):
        lines.append("    " + ", ".join(fmt % v for v in values[i:i + per_line]) + ",")
    lines.append("};")
    return "\n".join(lines)


def render(guard, banner):
    t = build_tables()
    owner = [edge_owner(e) for e in range(12)]
    face_c = [c for cyc, _ in FACES for c in cyc]
    edge_at = [255] * 24
    for e, (lo, axis) in enumerate(owner):
        edge_at[axis * 8 + lo[0] * 4 + lo[1] * 2 + lo[2]] = e
    corner_at = [0] * 8
    for i, c in enumerate(CORNERS):
        corner_at[int(c[0]) * 4 + int(c[1]) * 2 + int(c[2])] = i
    for config in range(256):  # the kernels rely on this: a cube without an ambiguous face never has a centre vertex
        if not ambiguous_faces(config):
            assert all(CENTRE not in tri for tri in tile_cube(config))
    parts = [
        f"/* {banner}",
        " * GENERATED by tools/gen_mc_tables.py -- do not edit.  See that script for the construction.",
        " * Corner i offsets (axis0, axis1, axis2); case index bit i set iff value(corner i) > level. */",
        f"#ifndef {guard}", f"#define {guard}", "#include <stdint.h>", "#ifndef HV_MC_TABLE",
        "#define HV_MC_TABLE static const", "#endif",
        f"#define HV_MC_NUM_VARIANTS {len(t['var_ntri'])}",
        f"#define HV_MC_NUM_TRIS {len(t['tris'])}",
        f"#define HV_MC_MAX_TRIS_PER_CUBE {max(t['var_ntri'])}",
        _c_array("uint8_t", "kMcCornerOff", [int(v) for c in CORNERS for v in c], 12),
        "/* corner id from its offsets: [d0*4 + d1*2 + d2] */",
        _c_array("uint8_t", "kMcCornerAt", corner_at, 8),
        "/* edge e joins corners kMcEdgeCorner[2e] (lower) and [2e+1] (upper); owner voxel offset + axis below */",
        _c_array("uint8_t", "kMcEdgeCorner", [c for e in range(12) for c in sorted(EDGES[e], key=lambda k: tuple(CORNERS[k]))], 12),
        _c_array("uint8_t", "kMcEdgeOwner", [v for (lo, axis) in owner for v in (*lo, axis)], 16),
        "/* edge id from (axis, offset of its lower corner inside the cube): [axis*8 + d0*4 + d1*2 + d2], 255 = none */",
        _c_array("uint8_t", "kMcEdgeAt", edge_at, 8),
        "/* face f: its four corners in cyclic order */",
        _c_array("uint8_t", "kMcFaceCorner", face_c, 12),
        _c_array("uint8_t", "kMcAmbCount", t["amb_count"]),
        "/* ambiguous faces of a configuration, 3 bits each, in test order */",
        _c_array("uint32_t", "kMcAmbFaces", t["amb_faces"], 8, "0x%xu"),
        _c_array("uint16_t", "kMcVariantBase", t["var_base"]),
        _c_array("uint8_t", "kMcVarNtri", t["var_ntri"]),
        _c_array("uint16_t", "kMcVarTriOff", t["var_off"]),
        "/* per variant: how many triangle corners reference edge e (4 bits per edge; e = 12 is the centre vertex) */",
        _c_array("uint64_t", "kMcVarMult", t["var_mult"], 4, "0x%xull"),
        _c_array("uint8_t", "kMcTris", [e for tri in t["tris"] for e in tri], 24),
        "#endif", "",
    ]
    return "\n".join(parts)


def main():
    outs = [
        (os.path.join(ROOT, "hive_b200", "csrc", "hv_mc_tables.h"), "HV_MC_TABLES_H",
         "Marching-cubes tables of the hive_b200 CUDA kernels."),
        (os.path.join(ROOT, "oracle", "mc_tables.h"), "HVO_MC_TABLES_H",
         "ORACLE (test infrastructure) -- marching-cubes tables of oracle/mc_oracle.c."),
    ]
    for path, guard, banner in outs:
        with open(path, "w") as f:
            f.write(render(guard, banner))
        print("wrote", path)


if __name__ == "__main__":
    sys.exit(main())
