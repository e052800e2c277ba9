#!/usr/bin/env python
"""Generate the marching-cubes tables of the CUDA kernels (hive_b200/csrc/hv_mc_tables.h).

    python tools/gen_mc_tables.py          # rewrites hive_b200/csrc/hv_mc_tables.h

The C oracle (oracle/mc_oracle.c) does NOT read this file or its output: it derives every cube's tiling at run time
with its own, separately written code, from the specification in DESIGN.md section 2.1 ("Marching-cubes
specification").  tests/test_mc_tables.py compares the two derivations for every (configuration, face resolution,
tunnel) triple.

Why generated: the reference extracts its mesh with ``skimage.measure.marching_cubes`` (Lewiner et al. 2003),
scikit-image==0.19.1, a third-party dependency absent from /root/reference and from this image (SURVEY.md §8(c),
App. C) -- its look-up tables cannot be consulted, and they are not typed from memory.  What is published is the
algorithm, and this script follows it step by step:

  1. corner / edge numbering and the case index follow the skimage Cython (App. C): bit i set iff v_i > level;
  2. the configuration is classified into Lewiner's base cases 0..14 (by the adjacency structure of the minority
     corners) and marked *inverted* when the positive corners are the majority;
  3. each ambiguous face is resolved with the face test ("asymptotic decider", A*C - B*D); Lewiner's test result is
     "the minority corners are joined across the face", here kept as the bit p = "the POSITIVE corners are joined";
  4. the sub-case follows Lewiner's decision tree: 3.1/3.2, 4.1.1/4.1.2, 6.1.1/6.1.2/6.2, 7.1/7.2/7.3/7.4.1/7.4.2,
     10.1.1/10.1.1_/10.1.2/10.2, 12.1.1/12.1.1_/12.1.2/12.2, 13.1/13.2/13.3/13.4/13.5.1/13.5.2 and their inverses;
     the *interior test* (the trilinear interpolant joins two diagonally opposite components through the cube: a
     "tunnel") is evaluated exactly where Lewiner's ``process_cube`` evaluates it -- not, e.g., for 10.1.1_ / 12.1.1_;
  5. the surface inside the cube is bounded by the iso-contours on the six faces; they chain into closed loops; every
     loop bounds a disk, except that a tunnel sub-case joins two loops by a tube; sub-cases 7.3, 10.2, 12.2, 13.3 and
     13.4 fan their 8-, 9- or 12-gon around an extra centre vertex and 6.1.2 puts the centre vertex into its tube, as
     Lewiner's tilings do -- which reproduces his per-sub-case triangle counts (SURVEY.md App. C inventory);
  6. orientation: right-hand normals point towards increasing values in the (axis0, axis1, axis2) output frame, which
     is what skimage returns for gradient_direction='descent' (its default; App. C).

What cannot be reproduced without Lewiner's tables (DESIGN.md section 2.1 lists these as open): which diagonals
triangulate a polygon inside one cube, and the reference edge of the interior test for cases 6, 7 and 13 where the
geometry leaves two or three equivalent candidates (case 12 has a unique one; cases 4 and 10 use no edge).
"""
from __future__ import annotations

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


def edge_axis(e):
    return edge_owner(e)[1]


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
NEIGHBOURS = {i: [j for j in range(8) if np.abs(CORNERS[i] - CORNERS[j]).sum() == 1] for i in range(8)}
OPPOSITE = {i: [j for j in range(8) if np.abs(CORNERS[i] - CORNERS[j]).sum() == 3][0] for i in range(8)}


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
    """Closed, oriented contour loops (lists of edge indices, each starting at its lowest edge, sorted by that edge)."""
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


# ---- Lewiner's classification -------------------------------------------------------------------------------------------
def classify(config):
    """(base case 0..14, inverted) of a sign configuration.  `inverted`: the positive corners are the majority, i.e.
    the tests of Lewiner's tables carry a negative sign for it."""
    pos = [i for i in range(8) if (config >> i) & 1]
    inv = len(pos) > 4
    m = [i for i in range(8) if i not in pos] if inv else pos
    deg = {i: sum(1 for j in NEIGHBOURS[i] if j in m) for i in m}
    n_adj = sum(deg.values()) // 2
    if len(m) == 0:
        case = 0
    elif len(m) == 1:
        case = 1
    elif len(m) == 2:
        d = int(np.abs(CORNERS[m[0]] - CORNERS[m[1]]).sum())
        case = {1: 2, 2: 3, 3: 4}[d]
    elif len(m) == 3:
        case = {2: 5, 1: 6, 0: 7}[n_adj]
    else:
        if n_adj == 4:
            case = 8
        elif n_adj == 3:
            if max(deg.values()) == 3:
                case = 9
            else:  # a path of four corners: the two chiral zig-zags 11 / 14 (neither has an ambiguous face)
                ends = [i for i in m if deg[i] == 1]
                mids = [i for i in m if deg[i] == 2]
                a, b = (mids if tuple(sorted((ends[0], mids[0]))) in [tuple(sorted(e)) for e in EDGES] else mids[::-1])
                chir = np.linalg.det(np.array([CORNERS[a] - CORNERS[ends[0]], CORNERS[b] - CORNERS[a],
                                               CORNERS[ends[1]] - CORNERS[b]], dtype=float))
                case = 11 if chir > 0 else 14
        elif n_adj == 2:
            case = 12 if max(deg.values()) == 2 else 10
        else:
            assert n_adj == 0
            case = 13
    return case, inv


def _common_face_edges(f, corner):
    """Edges of face f incident to `corner`."""
    cyc = FACES[f][0]
    k = cyc.index(corner)
    return [EDGE_OF[frozenset((corner, cyc[(k + 1) % 4]))], EDGE_OF[frozenset((corner, cyc[(k - 1) % 4]))]]


def _pick_edge(cands):
    """Convention where the geometry leaves several equivalent reference edges: the one along the lowest output axis
    (axis 0 = Lewiner's z first, then axis 1 = y, then axis 2 = x); ties cannot occur among edges at one corner."""
    return min(cands, key=lambda e: (edge_axis(e), e))


def _parallel_edges(e):
    """The reference edge and its three parallel edges in the layout of Lewiner's test_interior: (A, B, C, D) with A the
    reference edge, C the diagonally opposite one, B and D the two adjacent ones; each as (from, to) corners, all
    pointing the way EDGES[e] points."""
    a0, a1 = EDGES[e]
    axis = edge_axis(e)
    direction = CORNERS[a1] - CORNERS[a0]
    out = [(a0, a1)]
    others = []
    for c0 in range(8):
        if c0 == a0 or CORNERS[c0][axis] != CORNERS[a0][axis]:
            continue
        c1 = [j for j in range(8) if (CORNERS[j] == CORNERS[c0] + direction).all()][0]
        others.append((int(np.abs(CORNERS[c0] - CORNERS[a0]).sum()), c0, c1))
    adj = sorted((c0, c1) for d, c0, c1 in others if d == 1)
    opp = [(c0, c1) for d, c0, c1 in others if d == 2]
    assert len(adj) == 2 and len(opp) == 1
    return [out[0], adj[0], opp[0], adj[1]]


SLAB_EDGES = [(0, 4), (3, 7), (2, 6), (1, 5)]  # A, B, C, D of Lewiner's test_interior for cases 4 and 10


def subcase(config, p_bits):
    """Lewiner's sub-case for a configuration whose ambiguous faces are resolved by `p_bits` (bit k: the positive
    corners of the k-th ambiguous face are joined).  Returns (label, interior) where `interior` is None or a dict
    {kind: 'slab' | 'edge', edges: [(from, to)] * 4, inverted, tunnel_if_test} describing the interior test that
    decides between label + '.1' (two disks) and label + '.2' (a tube)."""
    case, inv = classify(config)
    amb = ambiguous_faces(config)
    # Lewiner's face-test result: the MINORITY corners are joined across the face
    T = [bool((p_bits >> k) & 1) != inv for k in range(len(amb))]
    pos = [i for i in range(8) if (config >> i) & 1]
    m = [i for i in range(8) if i not in pos] if inv else pos

    def edge_test(e, tunnel_if_test=False):
        return dict(kind="edge", edges=_parallel_edges(e), inverted=inv, tunnel_if_test=tunnel_if_test)

    if case in (0, 1, 2, 5, 8, 9, 11, 14):
        assert not amb
        return str(case), None
    if case == 3:
        return ("3.2" if T[0] else "3.1"), None
    if case == 4:
        assert not amb
        return "4.1", dict(kind="slab", edges=SLAB_EDGES, inverted=inv, tunnel_if_test=False)
    if case == 6:
        if T[0]:
            return "6.2", None
        iso = [i for i in m if not any(j in m for j in NEIGHBOURS[i])][0]
        return "6.1", edge_test(_pick_edge(_common_face_edges(amb[0], iso)))
    if case == 7:
        n = sum(T)
        if n < 3:
            return ("7.1", "7.2", "7.3")[n], None
        c = [i for i in range(8) if i not in m and all(j in m for j in NEIGHBOURS[i])][0]
        edges_c = [EDGE_OF[frozenset((c, j))] for j in NEIGHBOURS[c]]
        return "7.4", edge_test(_pick_edge(edges_c), tunnel_if_test=True)
    if case in (10, 12):
        assert not inv and len(amb) == 2
        if T[0] and T[1]:
            return f"{case}.1.1_", None
        if T[0] or T[1]:
            return f"{case}.2", None
        if case == 10:
            return "10.1", dict(kind="slab", edges=SLAB_EDGES, inverted=False, tunnel_if_test=False)
        common = set(EDGE_OF[frozenset((FACES[amb[0]][0][k], FACES[amb[0]][0][(k + 1) % 4]))] for k in range(4)) & \
            set(EDGE_OF[frozenset((FACES[amb[1]][0][k], FACES[amb[1]][0][(k + 1) % 4]))] for k in range(4))
        assert len(common) == 1
        return "12.1", edge_test(common.pop())
    assert case == 13 and not inv and len(amb) == 6
    true_faces = [amb[k] for k in range(6) if T[k]]
    false_faces = [amb[k] for k in range(6) if not T[k]]

    def share_corner(faces):
        common = set(range(8))
        for f in faces:
            common &= set(FACES[f][0])
        return common

    n = len(true_faces)
    if n == 0:
        return "13.1", None
    if n == 1:
        return "13.2", None
    if n == 2:
        return ("13.3" if not _opposite_faces(*true_faces) else "13.x"), None
    if n == 3:
        common = share_corner(true_faces)
        if len(common) != 1:
            return "13.x", None
        c = common.pop()
        if c in pos:
            return "13.4", None
        p_star = OPPOSITE[c]
        edges_p = [EDGE_OF[frozenset((p_star, j))] for j in NEIGHBOURS[p_star]]
        return "13.5", edge_test(_pick_edge(edges_p))
    if n == 4:
        return ("13.3_" if not _opposite_faces(*false_faces) else "13.x"), None
    if n == 5:
        return "13.2_", None
    return "13.1_", None


def _opposite_faces(f, g):
    return not (set(FACES[f][0]) & set(FACES[g][0]))


# Triangles per sub-case as SURVEY.md App. C records them from Lewiner's table widths (the only reference-derived
# numbers available on this box); "13.x" = face-test combinations that cannot occur in exact arithmetic.
LEWINER_NTRI = {
    "1": 1, "2": 2, "3.1": 2, "3.2": 4, "4.1.1": 2, "4.1.2": 6, "5": 3, "6.1.1": 3, "6.1.2": 9, "6.2": 5,
    "7.1": 3, "7.2": 5, "7.3": 9, "7.4.1": 5, "7.4.2": 9, "8": 2, "9": 4, "10.1.1": 4, "10.1.1_": 4, "10.1.2": 8,
    "10.2": 8, "11": 4, "12.1.1": 4, "12.1.1_": 4, "12.1.2": 8, "12.2": 8, "13.1": 4, "13.1_": 4, "13.2": 6,
    "13.2_": 6, "13.3": 10, "13.3_": 10, "13.4": 12, "13.5.1": 6, "13.5.2": 10, "14": 4,
}
CENTRE_FAN = ("7.3", "10.2", "12.2", "13.3", "13.3_", "13.4")  # the largest loop is fanned around the centre vertex


# ---- triangulation rules ------------------------------------------------------------------------------------------------
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


def _fan(loop):
    return tuple((CENTRE, loop[i], loop[(i + 1) % len(loop)]) for i in range(len(loop)))


def _same_face(*edges):
    """All the given edge vertices lie on one cube face."""
    common = set(range(6))
    for e in edges:
        common &= _FACES_OF_EDGE[e]
    return bool(common)


def _tube(la, lb):
    """Tube between loops A and B (both keep their orientation: the surface lies to the left of either) as a zipper
    without interior vertex: len(A) + len(B) triangles, each with one contour segment and its apex on the other loop.
    A *bridge* joins a vertex of A to a vertex of B; a bridge lying in a cube face is avoided where possible (7.4.2
    admits no tube without three of them): budgets 0, 1, 2, ... of in-face bridges are tried in turn, and for a budget
    the first zipper in the order "start s in A, start t in B, advance on A before advancing on B" wins.  No triangle
    may have its three vertices on one cube face, and neither loop may be consumed in one uninterrupted run (that
    would be a cone, not a tube).  Returns (triangles, number of in-face bridges)."""
    na, nb = len(la), len(lb)
    for budget in range(na + nb + 1):
        for s in range(na):
            for t in range(nb):
                c0 = int(_coplanar(la[s], lb[t]))
                if c0 > budget:
                    continue
                out = []

                def rec(i, j, used, a_at, a_split, b_at, b_split):
                    # i segments of A and j segments of B consumed; the current bridge joins A[s+i] and B[t-j].
                    # a_at: the j at which A advanced first, a_split: A has advanced at two different j (else the
                    # segments of A would form a cone over one vertex of B and use its bridge twice); same for B.
                    if i == na and j == nb:
                        return a_split and b_split
                    a, b = la[(s + i) % na], lb[(t - j) % nb]
                    if i < na:
                        a2 = la[(s + i + 1) % na]
                        cost = 0 if (i + 1 == na and j == nb) else int(_coplanar(a2, b))
                        if used + cost <= budget and not _same_face(a, a2, b):
                            out.append((a, a2, b))
                            if rec(i + 1, j, used + cost, j if a_at < 0 else a_at, a_split or (a_at >= 0 and a_at != j),
                                   b_at, b_split):
                                return True
                            out.pop()
                    if j < nb:
                        b2 = lb[(t - j - 1) % nb]
                        cost = 0 if (i == na and j + 1 == nb) else int(_coplanar(a, b2))
                        if used + cost <= budget and not _same_face(b2, b, a):
                            out.append((b2, b, a))
                            if rec(i, j + 1, used + cost, a_at, a_split, i if b_at < 0 else b_at,
                                   b_split or (b_at >= 0 and b_at != i)):
                                return True
                            out.pop()
                    return False

                if rec(0, 0, c0, -1, False, -1, False):
                    return tuple(out), budget
    return None, -1


def _tube_centre(la, lb):
    """Tube with the centre vertex (Lewiner's 6.1.2): the segments A[k] -> A[k+1] and B[j] -> B[j+1] close a quad with
    the bridges A[k+1]-B[j] and A[k]-B[j+1]; everything else is fanned around the centre.  len(A) + len(B) + 2
    triangles.  First valid (k, j, quad diagonal) in lexicographic order."""
    na, nb = len(la), len(lb)
    for k in range(na):
        for j in range(nb):
            a0, a1, b0, b1 = la[k], la[(k + 1) % na], lb[j], lb[(j + 1) % nb]
            if _coplanar(a1, b0) or _coplanar(a0, b1):
                continue
            for diag in (0, 1):
                if diag == 0 and not _coplanar(a0, b0):
                    quad = ((a0, a1, b0), (a0, b0, b1))
                elif diag == 1 and not _coplanar(a1, b1):
                    quad = ((a0, a1, b1), (a1, b0, b1))
                else:
                    continue
                ring = [la[(k + 1 + i) % na] for i in range(na)] + [lb[(j + 1 + i) % nb] for i in range(nb)]
                return quad + _fan(ring)
    return None


def tile_cube(config, p_bits=0, tunnel=False):
    """Triangles (edge-index triples, 12 = centre vertex) of one cube and the sub-case label."""
    label, interior = subcase(config, p_bits)
    assert not (tunnel and interior is None)
    if interior is not None:
        label += ".2" if tunnel else ".1"
    loops = cube_loops(config, p_bits)
    tris = []
    if tunnel:
        if label == "13.5.2":  # three loops: the triangle around the lone positive corner joins the hexagon
            hexa = [l for l in loops if len(l) == 6]
            assert len(hexa) == 1 and len(loops) == 3
            p_star = interior["edges"][0][0] if (config >> interior["edges"][0][0]) & 1 else interior["edges"][0][1]
            ring = sorted(EDGE_OF[frozenset((p_star, j))] for j in NEIGHBOURS[p_star])
            small = [l for l in loops if sorted(l) == ring]
            assert len(small) == 1
            rest = [l for l in loops if l is not hexa[0] and l is not small[0]]
            pair = (hexa[0], small[0])
        else:
            assert len(loops) == 2
            pair, rest = (loops[0], loops[1]), []
        t = _tube_centre(*pair) if label == "6.1.2" else _tube(*pair)[0]
        assert t is not None, f"no tube for configuration {config} ({label})"
        tris.extend(t)
        loops = rest
    big = max(range(len(loops)), key=lambda i: (len(loops[i]), -i)) if loops else -1
    for i, loop in enumerate(loops):
        if label in CENTRE_FAN and i == big:
            t = _fan(loop)
        else:
            t = _triangulate(loop)
            if t is None:  # only the face-test combinations of case 13 that cannot occur in exact arithmetic
                assert label == "13.x" and all(CENTRE not in tri for tri in tris)
                t = _fan(loop)
        tris.extend(t)
    if label in LEWINER_NTRI:
        assert len(tris) == LEWINER_NTRI[label], (config, p_bits, tunnel, label, len(tris))
    return tris, label


FLT_EPSILON = float(np.finfo(np.float32).eps)


def face_bit(config, f, v):
    """Lewiner's test_face as the bit p = "the positive corners of ambiguous face f are joined" (step F)."""
    cyc = FACES[f][0]
    k = 0 if (config >> cyc[0]) & 1 else 1
    pos = float(v[cyc[k]]) * float(v[cyc[k + 2]])
    neg = float(v[cyc[k + 1]]) * float(v[cyc[(k + 3) % 4]])
    return pos - neg > -FLT_EPSILON


def interior_tunnel(interior, v):
    """Lewiner's test_interior on the descriptor returned by subcase() (step I); True = use the tube tiling."""
    v = [float(x) for x in v]
    ret_pos, ret_neg = (not interior["inverted"]), interior["inverted"]
    (a0, a1), (b0, b1), (c0, c1), (d0, d1) = [(v[p], v[q]) for p, q in interior["edges"]]
    with np.errstate(all="ignore"):
        if interior["kind"] == "slab":
            dA, dB, dC, dD = a1 - a0, b1 - b0, c1 - c0, d1 - d0
            qa = np.float64(dA * dC - dB * dD)
            qb = np.float64(c0 * dA + a0 * dC - d0 * dB - b0 * dD)
            t = float(-qb / (2 * qa))
            if t < 0 or t > 1:
                result = ret_pos
                return result if interior["tunnel_if_test"] else not result
            At, Bt, Ct, Dt = a0 + dA * t, b0 + dB * t, c0 + dC * t, d0 + dD * t
        else:
            t = float(np.float64(a0) / np.float64(a0 - a1))
            At, Bt, Ct, Dt = 0.0, b0 + (b1 - b0) * t, c0 + (c1 - c0) * t, d0 + (d1 - d0) * t
    test = (1 if At >= 0 else 0) | (2 if Bt >= 0 else 0) | (4 if Ct >= 0 else 0) | (8 if Dt >= 0 else 0)
    if test == 5:
        result = ret_pos if At * Ct - Bt * Dt < FLT_EPSILON else ret_neg
    elif test == 10:
        result = ret_pos if At * Ct - Bt * Dt >= FLT_EPSILON else ret_neg
    elif test in (7, 11, 13, 14, 15):
        result = ret_neg
    else:
        result = ret_pos
    return result if interior["tunnel_if_test"] else not result


def decide(v):
    """(configuration, p_bits, tunnel, label) of a cube with corner values v (level subtracted)."""
    config = sum(1 << i for i in range(8) if v[i] > 0)
    if config in (0, 255):
        return config, 0, False, "0"
    amb = ambiguous_faces(config)
    p_bits = sum(1 << k for k, f in enumerate(amb) if face_bit(config, f, v))
    label, interior = subcase(config, p_bits)
    tunnel = interior is not None and interior_tunnel(interior, v)
    if interior is not None:
        label += ".2" if tunnel else ".1"
    return config, p_bits, tunnel, label


def interior_word(interior):
    """Descriptor of one interior test: bits 0-1 kind (1 slab, 2 edge), bit 2 inverted, bit 3 tunnel_if_test, then
    the eight corners A0 A1 B0 B1 C0 C1 D0 D1, three bits each."""
    w = {"slab": 1, "edge": 2}[interior["kind"]] | (4 if interior["inverted"] else 0) | \
        (8 if interior["tunnel_if_test"] else 0)
    shift = 4
    for c0, c1 in interior["edges"]:
        w |= (c0 << shift) | (c1 << (shift + 3))
        shift += 6
    return w


def build_tables():
    amb_count, amb_faces, var_base, int_base, int_mask = [], [], [], [], []
    var_ntri, var_off, var_mult, tris, int_words, labels = [], [], [], [], [], []

    def add_variant(t, label):
        var_ntri.append(len(t))
        var_off.append(len(tris))
        mult = [0] * 13
        for tri in t:
            assert len(set(tri)) == 3
            for e in tri:
                mult[e] += 1
        assert max(mult) < 16
        var_mult.append(sum(m << (4 * e) for e, m in enumerate(mult)))
        tris.extend(t)
        labels.append(label)

    for config in range(256):
        amb = ambiguous_faces(config)
        amb_count.append(len(amb))
        amb_faces.append(sum(f << (3 * k) for k, f in enumerate(amb)))
        var_base.append(len(var_ntri))
        if config in (0, 255):
            add_variant([], "0")
            int_base.append(255)
            int_mask.append(0)
            continue
        tunnels = []
        for bits in range(1 << len(amb)):
            _, interior = subcase(config, bits)
            add_variant(*tile_cube(config, bits, False))
            if interior is not None:
                tunnels.append((bits, interior))
        int_base.append(len(int_words) if tunnels else 255)
        int_mask.append(sum(1 << bits for bits, _ in tunnels))
        for bits, interior in tunnels:  # the tunnel tilings follow the 2^namb plain ones, in p_bits order
            add_variant(*tile_cube(config, bits, True))
            int_words.append(interior_word(interior))
    assert len(int_words) < 255
    return dict(amb_count=amb_count, amb_faces=amb_faces, var_base=var_base, var_ntri=var_ntri, var_off=var_off,
                var_mult=var_mult, tris=tris, int_base=int_base, int_mask=int_mask, int_words=int_words, labels=labels)


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
            for tunnel in ((False, True) if subcase(config, 0)[1] else (False,)):
                assert all(CENTRE not in tri for tri in tile_cube(config, 0, tunnel)[0])
    parts = [
        f"/* {banner}",
        " * GENERATED by tools/gen_mc_tables.py -- do not edit.  See that script for the construction.",
        " * Corner i offsets (axis0, axis1, axis2); case index bit i set iff value(corner i) > level. */",
        f"#ifndef {guard}", f"#define {guard}", "#include <stdint.h>", "#ifndef HV_MC_TABLE",
        "#define HV_MC_TABLE static const", "#endif",
        f"#define HV_MC_NUM_VARIANTS {len(t['var_ntri'])}",
        f"#define HV_MC_NUM_TRIS {len(t['tris'])}",
        f"#define HV_MC_MAX_TRIS_PER_CUBE {max(t['var_ntri'])}",
        f"#define HV_MC_NUM_INTERIOR {len(t['int_words'])}",
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
        "/* tiling of (configuration, p_bits) = kMcVariantBase[configuration] + p_bits, bit k of p_bits: the positive",
        " * corners of the k-th ambiguous face are joined; tunnel tilings follow at + 2^namb + rank of p_bits in the mask */",
        _c_array("uint16_t", "kMcVariantBase", t["var_base"]),
        "/* interior test: which p_bits need one (bit mask), and where their descriptors start (255 = never) */",
        _c_array("uint64_t", "kMcInteriorMask", t["int_mask"], 4, "0x%xull"),
        _c_array("uint8_t", "kMcInteriorBase", t["int_base"]),
        "/* descriptor: bits 0-1 kind (1 = slab test of cases 4/10, 2 = reference-edge test), bit 2 = inverted (Lewiner's",
        " * s < 0), bit 3 = the tube is chosen when the test is TRUE (case 7), then corners A0 A1 B0 B1 C0 C1 D0 D1 (3 bits each) */",
        _c_array("uint32_t", "kMcInterior", t["int_words"], 8, "0x%xu"),
        _c_array("uint8_t", "kMcVarNtri", t["var_ntri"]),
        _c_array("uint16_t", "kMcVarTriOff", t["var_off"]),
        "/* per variant: how many triangle corners reference edge e (4 bits per edge; e = 12 is the centre vertex) */",
        _c_array("uint64_t", "kMcVarMult", t["var_mult"], 4, "0x%xull"),
        _c_array("uint8_t", "kMcTris", [e for tri in t["tris"] for e in tri], 24),
        "#endif", "",
    ]
    return "\n".join(parts)


def main():
    path = os.path.join(ROOT, "hive_b200", "csrc", "hv_mc_tables.h")
    with open(path, "w") as f:
        f.write(render("HV_MC_TABLES_H", "Marching-cubes tables of the hive_b200 CUDA kernels."))
    print("wrote", path)


if __name__ == "__main__":
    sys.exit(main())
