/* ORACLE -- test infrastructure, not product code.
 *
 * Serial CPU restatement of the marching-cubes step of TSDFVolume.get_mesh
 * (third_party/tsdf_fusion_python/fusion.py:338, `measure.marching_cubes(tsdf_vol, level=0)`).
 *
 * PARITY UNPINNED.  The arithmetic lives in scikit-image==0.19.1 (requirements.txt:18,
 * skimage/measure/_marching_cubes_lewiner_cy.pyx), a third-party dependency that is neither vendored under
 * /root/reference nor installed on the build or GPU box, and the reference has no test or golden vector for it
 * (SURVEY.md §4, §8(c)).  This file restates the published algorithm (Lewiner, Lopes, Vieira, Tavares 2003) and the
 * wrapper contract recorded in SURVEY.md App. C:
 *   - cubes are visited with axis 0 outermost and axis 2 innermost, cubes 0..dim-2 per axis;
 *   - case index bit i is set iff corner i (Lewiner numbering) is > level, values taken as double(v) - level;
 *   - one vertex per sign-changing lattice edge, shared between cubes through two sliding layers of edge->index maps
 *     (numbered in first-touch order), placed at the centre of mass of the two corners weighted by
 *     1 / (FLT_EPSILON + |v - level|), float64 arithmetic, stored as float32;
 *   - normals: in-cube finite differences pointing down the gradient, interpolated with the same weights, added once
 *     per triangle corner that references the vertex (float32 accumulation in visiting order), normalised at the end;
 *   - faces come out oriented as skimage's default gradient_direction='descent' leaves them (right-hand normal
 *     towards increasing values in the (axis0, axis1, axis2) frame).
 *
 * INDEPENDENT OF THE PRODUCT'S TABLES.  Lewiner's look-up tables cannot be consulted on this box, so both this file
 * and the CUDA kernels follow the written specification in DESIGN.md section 2.1: the kernels through tables generated
 * offline by tools/gen_mc_tables.py (Python), this file by deriving each cube's tiling at run time from the cube
 * geometry with the code below (C, written separately; results are cached per (configuration, face resolution,
 * tunnel) because the bench times this file at 512^3).  Nothing here includes or reads hv_mc_tables.h.
 * Specification steps, in the order of the code:
 *   G  geometry: corners, edges, faces (counter-clockwise seen from outside), which faces an edge lies on;
 *   F  face test: p = "the positive corners of an ambiguous face are joined" = (pos - neg > -FLT_EPSILON) with pos / neg
 *      the products of the positive / negative diagonal (Lewiner's test_face incl. its epsilon band);
 *   C  base case 0..14 from the adjacency of the minority corners; inverted = positives are the majority;
 *   S  sub-case by Lewiner's decision tree on T = p xor inverted ("the minority is joined");
 *   I  interior test (Lewiner's test_interior, slab form for cases 4 / 10, reference-edge form for 6 / 7 / 12 / 13),
 *      evaluated only where his process_cube evaluates it;
 *   L  contour loops: segments directed with the positive corners on their left seen from outside, loops started at
 *      their lowest edge and listed by that edge;
 *   T  triangulation: disk = first triangulation without in-face chords (recursion on (i, j), apex k ascending);
 *      centre fan for 7.3 / 10.2 / 12.2 / 13.3 / 13.4; tube = zipper with the smallest number of in-face bridges;
 *      6.1.2 = tube with the centre vertex.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- G: geometry ---------------------------------------------------------------------------------------------------- */
enum { NC = 8, NE = 12, NF = 6, CENTRE = 12, MAX_TRI = 16 };

static int g_ready = 0;
static int g_corner[NC][3];     /* offsets along (axis0, axis1, axis2) */
static int g_edge[NE][2];       /* Lewiner's edge order and direction */
static int g_edge_axis[NE];
static unsigned g_edge_faces[NE]; /* bit (2 * axis + side) for every face the edge lies on */
static int g_edge_id[NC][NC];
static int g_face[NF][4];       /* face 2 * axis + side: corners counter-clockwise seen from outside */
static int g_nbr[NC][3];
static int g_opposite[NC];

static void geometry_init(void) {
    if (g_ready) return;
    for (int i = 0; i < NC; ++i) { /* Lewiner / skimage: v0(0,0,0) v1(1,0,0) v2(1,1,0) v3(0,1,0), v4..v7 the same at z = 1 */
        const int y = (i >> 1) & 1, x = (i & 1) ^ y, z = (i >> 2) & 1;
        g_corner[i][0] = z; /* skimage's x is array axis 2, its z is array axis 0 */
        g_corner[i][1] = y;
        g_corner[i][2] = x;
    }
    for (int i = 0; i < 4; ++i) {
        g_edge[i][0] = i, g_edge[i][1] = (i + 1) & 3;
        g_edge[4 + i][0] = 4 + i, g_edge[4 + i][1] = 4 + ((i + 1) & 3);
        g_edge[8 + i][0] = i, g_edge[8 + i][1] = i + 4;
    }
    memset(g_edge_id, 0xff, sizeof(g_edge_id));
    for (int e = 0; e < NE; ++e) {
        const int a = g_edge[e][0], b = g_edge[e][1];
        g_edge_id[a][b] = g_edge_id[b][a] = e;
        g_edge_faces[e] = 0;
        for (int k = 0; k < 3; ++k) {
            if (g_corner[a][k] != g_corner[b][k]) g_edge_axis[e] = k;
            else g_edge_faces[e] |= 1u << (2 * k + g_corner[a][k]);
        }
    }
    for (int axis = 0; axis < 3; ++axis)
        for (int side = 0; side < 2; ++side) {
            int u = (axis + 1) % 3, v = (axis + 2) % 3; /* (u, v, +axis) is right-handed */
            if (!side) { const int t = u; u = v; v = t; }
            static const int ring[4][2] = {{0, 0}, {1, 0}, {1, 1}, {0, 1}};
            for (int k = 0; k < 4; ++k)
                for (int i = 0; i < NC; ++i)
                    if (g_corner[i][axis] == side && g_corner[i][u] == ring[k][0] && g_corner[i][v] == ring[k][1])
                        g_face[2 * axis + side][k] = i;
        }
    for (int i = 0; i < NC; ++i) {
        int n = 0;
        for (int j = 0; j < NC; ++j) {
            const int d = abs(g_corner[i][0] - g_corner[j][0]) + abs(g_corner[i][1] - g_corner[j][1]) +
                          abs(g_corner[i][2] - g_corner[j][2]);
            if (d == 1) g_nbr[i][n++] = j;
            if (d == 3) g_opposite[i] = j;
        }
    }
    g_ready = 1;
}

static int in_face(int a, int b) { return (g_edge_faces[a] & g_edge_faces[b]) != 0; }
static int in_face3(int a, int b, int c) { return (g_edge_faces[a] & g_edge_faces[b] & g_edge_faces[c]) != 0; }
static int bit(int config, int i) { return (config >> i) & 1; }

static int face_ambiguous(int config, int f) {
    const int *c = g_face[f];
    return bit(config, c[0]) == bit(config, c[2]) && bit(config, c[1]) == bit(config, c[3]) &&
           bit(config, c[0]) != bit(config, c[1]);
}

/* ---- F: face test ---------------------------------------------------------------------------------------------------- */
static int face_positives_joined(int config, int f, const double v[8]) {
    const int *c = g_face[f];
    const int k = bit(config, c[0]) ? 0 : 1; /* c[k], c[k + 2]: the positive diagonal */
    const double pos = v[c[k]] * v[c[k + 2]], neg = v[c[k + 1]] * v[c[(k + 3) & 3]];
    return pos - neg > -(double)FLT_EPSILON;
}

/* ---- C: base case ---------------------------------------------------------------------------------------------------- */
typedef struct {
    int base;          /* Lewiner's case 0..14 */
    int inverted;      /* positives are the majority */
    int minority[4], n_minority;
} cube_class;

static cube_class classify(int config) {
    cube_class k;
    int n_pos = 0;
    for (int i = 0; i < NC; ++i) n_pos += bit(config, i);
    k.inverted = n_pos > 4;
    k.n_minority = 0;
    for (int i = 0; i < NC; ++i)
        if (bit(config, i) != k.inverted) k.minority[k.n_minority++] = i;
    int degree[4] = {0, 0, 0, 0}, links = 0, max_degree = 0;
    for (int a = 0; a < k.n_minority; ++a)
        for (int b = 0; b < k.n_minority; ++b)
            if (a != b && g_edge_id[k.minority[a]][k.minority[b]] >= 0) ++degree[a], ++links;
    links /= 2;
    for (int a = 0; a < k.n_minority; ++a)
        if (degree[a] > max_degree) max_degree = degree[a];
    switch (k.n_minority) {
    case 0: k.base = 0; break;
    case 1: k.base = 1; break;
    case 2:
        if (links) k.base = 2;
        else k.base = (g_opposite[k.minority[0]] == k.minority[1]) ? 4 : 3;
        break;
    case 3: k.base = links == 2 ? 5 : (links == 1 ? 6 : 7); break;
    default:
        if (links == 4) k.base = 8;
        else if (links == 3) k.base = max_degree == 3 ? 9 : 11; /* 11 and its mirror image 14: no ambiguity, one label */
        else if (links == 2) k.base = max_degree == 2 ? 12 : 10;
        else k.base = 13;
    }
    return k;
}

/* ---- S + I: sub-case and interior-test description -------------------------------------------------------------------- */
typedef struct {
    char label[24];
    int interior;        /* 0 none, 1 slab form, 2 reference-edge form */
    int tunnel_if_test;  /* case 7: the tube is the outcome of a TRUE test */
    int inverted;
    int ref[4][2];       /* A, B, C, D: (from, to) corners */
    int lone;            /* 13.5: the positive corner whose triangle the tube joins to the hexagon */
} subcase_t;

static int lowest_axis_edge(const int *cands, int n) {
    int best = cands[0];
    for (int i = 1; i < n; ++i)
        if (g_edge_axis[cands[i]] < g_edge_axis[best] ||
            (g_edge_axis[cands[i]] == g_edge_axis[best] && cands[i] < best))
            best = cands[i];
    return best;
}

static void reference_edge(subcase_t *s, int e) {
    /* A = the edge, C = the parallel edge diagonally across, B / D = the two adjacent parallel edges (lower start
     * corner first); all four run the way Lewiner's edge e runs */
    const int a0 = g_edge[e][0], a1 = g_edge[e][1], axis = g_edge_axis[e];
    s->interior = 2;
    s->ref[0][0] = a0, s->ref[0][1] = a1;
    int n_adj = 0;
    for (int c0 = 0; c0 < NC; ++c0) {
        if (c0 == a0 || g_corner[c0][axis] != g_corner[a0][axis]) continue;
        int c1 = -1;
        for (int j = 0; j < NC; ++j) {
            int same = 1;
            for (int k = 0; k < 3; ++k)
                if (g_corner[j][k] != g_corner[c0][k] + (g_corner[a1][k] - g_corner[a0][k])) same = 0;
            if (same) c1 = j;
        }
        if (g_edge_id[c0][a0] >= 0) { /* adjacent */
            const int slot = n_adj++ ? 3 : 1;
            s->ref[slot][0] = c0, s->ref[slot][1] = c1;
        } else {
            s->ref[2][0] = c0, s->ref[2][1] = c1;
        }
    }
}

static void slab_edges(subcase_t *s) {
    static const int abcd[4][2] = {{0, 4}, {3, 7}, {2, 6}, {1, 5}};
    s->interior = 1;
    memcpy(s->ref, abcd, sizeof(abcd));
}

static int faces_opposite(int f, int g) { return (f >> 1) == (g >> 1); }

/* pmask: bit f set iff face f is ambiguous and its positive corners are joined */
static subcase_t subcase(int config, int pmask) {
    subcase_t s;
    memset(&s, 0, sizeof(s));
    s.lone = -1;
    const cube_class k = classify(config);
    s.inverted = k.inverted;
    int amb[NF], namb = 0, joined[NF], n_joined = 0; /* joined: Lewiner's test result, "the minority is joined" */
    for (int f = 0; f < NF; ++f)
        if (face_ambiguous(config, f)) {
            joined[namb] = ((pmask >> f) & 1) != k.inverted;
            n_joined += joined[namb];
            amb[namb++] = f;
        }
    switch (k.base) {
    case 3: strcpy(s.label, joined[0] ? "3.2" : "3.1"); break;
    case 4:
        strcpy(s.label, "4.1");
        slab_edges(&s);
        break;
    case 6:
        if (joined[0]) { strcpy(s.label, "6.2"); break; }
        strcpy(s.label, "6.1");
        for (int a = 0; a < 3; ++a) { /* the lone minority corner: reference edge = one of its two edges in the face */
            const int c = k.minority[a];
            int lone = 1;
            for (int b = 0; b < 3; ++b)
                if (b != a && g_edge_id[c][k.minority[b]] >= 0) lone = 0;
            if (!lone) continue;
            int cands[2], n = 0;
            for (int q = 0; q < 4; ++q)
                if (g_face[amb[0]][q] == c) {
                    cands[n++] = g_edge_id[c][g_face[amb[0]][(q + 1) & 3]];
                    cands[n++] = g_edge_id[c][g_face[amb[0]][(q + 3) & 3]];
                }
            reference_edge(&s, lowest_axis_edge(cands, n));
        }
        break;
    case 7:
        if (n_joined < 3) { sprintf(s.label, "7.%d", n_joined + 1); break; }
        strcpy(s.label, "7.4");
        s.tunnel_if_test = 1;
        for (int c = 0; c < NC; ++c) { /* the majority corner surrounded by the three minority corners */
            if (bit(config, c) != k.inverted) continue;
            int all = 1, cands[3];
            for (int q = 0; q < 3; ++q) {
                if (bit(config, g_nbr[c][q]) == k.inverted) all = 0;
                cands[q] = g_edge_id[c][g_nbr[c][q]];
            }
            if (all) reference_edge(&s, lowest_axis_edge(cands, 3));
        }
        break;
    case 10:
    case 12:
        if (n_joined == 2) sprintf(s.label, "%d.1.1_", k.base);
        else if (n_joined == 1) sprintf(s.label, "%d.2", k.base);
        else {
            sprintf(s.label, "%d.1", k.base);
            if (k.base == 10) slab_edges(&s);
            else
                for (int e = 0; e < NE; ++e) /* the one edge the two ambiguous faces share */
                    if ((g_edge_faces[e] >> amb[0]) & (g_edge_faces[e] >> amb[1]) & 1u) reference_edge(&s, e);
        }
        break;
    case 13: {
        int t[NF], nt = 0, u[NF], nu = 0;
        for (int i = 0; i < namb; ++i) {
            if (joined[i]) t[nt++] = amb[i];
            else u[nu++] = amb[i];
        }
        strcpy(s.label, "13.x"); /* combinations that exact arithmetic excludes */
        if (nt == 0) strcpy(s.label, "13.1");
        else if (nt == 1) strcpy(s.label, "13.2");
        else if (nt == 2 && !faces_opposite(t[0], t[1])) strcpy(s.label, "13.3");
        else if (nt == 4 && !faces_opposite(u[0], u[1])) strcpy(s.label, "13.3_");
        else if (nt == 5) strcpy(s.label, "13.2_");
        else if (nt == 6) strcpy(s.label, "13.1_");
        else if (nt == 3 && !faces_opposite(t[0], t[1]) && !faces_opposite(t[0], t[2]) && !faces_opposite(t[1], t[2])) {
            int corner = -1; /* the corner the three faces share */
            for (int c = 0; c < NC; ++c)
                if (g_corner[c][t[0] >> 1] == (t[0] & 1) && g_corner[c][t[1] >> 1] == (t[1] & 1) &&
                    g_corner[c][t[2] >> 1] == (t[2] & 1))
                    corner = c;
            if (bit(config, corner)) strcpy(s.label, "13.4");
            else {
                strcpy(s.label, "13.5");
                s.lone = g_opposite[corner];
                int cands[3];
                for (int q = 0; q < 3; ++q) cands[q] = g_edge_id[s.lone][g_nbr[s.lone][q]];
                reference_edge(&s, lowest_axis_edge(cands, 3));
            }
        }
        break;
    }
    default: sprintf(s.label, "%d", k.base == 14 ? 11 : k.base);
    }
    return s;
}

/* Lewiner's test_interior; returns 1 when the tube tiling is to be used. */
static int interior_says_tunnel(const subcase_t *s, const double v[8]) {
    double t, At, Bt, Ct, Dt;
    const int ret_pos = !s->inverted, ret_neg = s->inverted; /* "return s > 0", "return s < 0" */
    int result;
    const int(*r)[2] = s->ref;
    if (s->interior == 1) {
        const double dA = v[r[0][1]] - v[r[0][0]], dB = v[r[1][1]] - v[r[1][0]], dC = v[r[2][1]] - v[r[2][0]],
                     dD = v[r[3][1]] - v[r[3][0]];
        const double a = dA * dC - dB * dD;
        const double b = v[r[2][0]] * dA + v[r[0][0]] * dC - v[r[3][0]] * dB - v[r[1][0]] * dD;
        t = -b / (2 * a);
        if (t < 0 || t > 1) {
            result = ret_pos;
            goto done;
        }
        At = v[r[0][0]] + dA * t;
        Bt = v[r[1][0]] + dB * t;
        Ct = v[r[2][0]] + dC * t;
        Dt = v[r[3][0]] + dD * t;
    } else {
        t = v[r[0][0]] / (v[r[0][0]] - v[r[0][1]]);
        At = 0;
        Bt = v[r[1][0]] + (v[r[1][1]] - v[r[1][0]]) * t;
        Ct = v[r[2][0]] + (v[r[2][1]] - v[r[2][0]]) * t;
        Dt = v[r[3][0]] + (v[r[3][1]] - v[r[3][0]]) * t;
    }
    {
        const int test = (At >= 0 ? 1 : 0) | (Bt >= 0 ? 2 : 0) | (Ct >= 0 ? 4 : 0) | (Dt >= 0 ? 8 : 0);
        switch (test) {
        case 5: result = (At * Ct - Bt * Dt < (double)FLT_EPSILON) ? ret_pos : ret_neg; break;
        case 10: result = (At * Ct - Bt * Dt >= (double)FLT_EPSILON) ? ret_pos : ret_neg; break;
        case 7: case 11: case 13: case 14: case 15: result = ret_neg; break;
        default: result = ret_pos;
        }
    }
done:
    return s->tunnel_if_test ? result : !result;
}

/* ---- L: contour loops -------------------------------------------------------------------------------------------------- */
typedef struct {
    int n, len[4], e[4][12];
} loops_t;

/* Adds the directed segment that cuts the corners ring[k0 .. k0 + span - 1] off face f. */
static void add_segment(int *next, int f, int k0, int span, int group_positive) {
    const int *c = g_face[f];
    const int before = c[(k0 + 3) & 3], first = c[k0 & 3], last = c[(k0 + span - 1) & 3], after = c[(k0 + span) & 3];
    const int e_in = g_edge_id[before][first], e_out = g_edge_id[last][after];
    /* Walking the face boundary counter-clockwise (seen from outside) one meets e_in, the group, e_out; the segment
     * e_out -> e_in cuts the corner with the group on its left.  Positives go on the left. */
    if (group_positive) next[e_out] = e_in;
    else next[e_in] = e_out;
}

static loops_t contour_loops(int config, int pmask) {
    int next[NE];
    memset(next, 0xff, sizeof(next));
    for (int f = 0; f < NF; ++f) {
        const int *c = g_face[f];
        int s[4], n_pos = 0;
        for (int k = 0; k < 4; ++k) n_pos += (s[k] = bit(config, c[k]));
        if (n_pos == 0 || n_pos == 4) continue;
        if (n_pos == 1 || n_pos == 3) {
            const int want = n_pos == 1;
            for (int k = 0; k < 4; ++k)
                if (s[k] == want) add_segment(next, f, k, 1, want);
        } else if (s[0] == s[2]) { /* ambiguous: joined positives -> the two negative corners are cut off */
            const int cut = ((pmask >> f) & 1) ? 0 : 1;
            for (int k = 0; k < 4; ++k)
                if (s[k] == cut) add_segment(next, f, k, 1, cut);
        } else {
            for (int k = 0; k < 4; ++k)
                if (s[k] && s[(k + 1) & 3]) add_segment(next, f, k, 2, 1);
        }
    }
    loops_t L;
    L.n = 0;
    int seen[NE] = {0};
    for (int start = 0; start < NE; ++start) {
        if (next[start] < 0 || seen[start]) continue;
        int n = 0;
        for (int e = start; !seen[e]; e = next[e]) {
            seen[e] = 1;
            L.e[L.n][n++] = e;
        }
        L.len[L.n++] = n;
    }
    return L;
}

/* ---- T: triangulation --------------------------------------------------------------------------------------------------- */
typedef struct {
    int n;
    uint8_t t[MAX_TRI][3];
} tiling_t;

static void push(tiling_t *o, int a, int b, int c) {
    o->t[o->n][0] = (uint8_t)a, o->t[o->n][1] = (uint8_t)b, o->t[o->n][2] = (uint8_t)c;
    ++o->n;
}

/* Disk: triangulates loop[i..j] (as a polygon closed by the chord j -> i); apex k ascending, first success wins.
 * A chord is allowed only between vertices that share no cube face. */
static int disk(const int *loop, int i, int j, tiling_t *o) {
    if (j - i < 2) return 1;
    for (int k = i + 1; k < j; ++k) {
        if (k != i + 1 && in_face(loop[i], loop[k])) continue;
        if (k + 1 != j && in_face(loop[k], loop[j])) continue;
        const int mark = o->n;
        if (disk(loop, i, k, o)) {
            push(o, loop[i], loop[k], loop[j]);
            if (disk(loop, k, j, o)) return 1;
        }
        o->n = mark;
    }
    return 0;
}

static void fan(const int *loop, int n, tiling_t *o) {
    for (int i = 0; i < n; ++i) push(o, CENTRE, loop[i], loop[(i + 1) % n]);
}

typedef struct {
    const int *a, *b;
    int na, nb, s, t, budget;
} zip_t;

/* a_at: the j at which A advanced first (-1: not yet), a_split: A has advanced at two different j -- otherwise all of
 * A's segments would be a cone over one vertex of B, whose bridge would then be used twice; likewise for B. */
static int zip(const zip_t *z, int i, int j, int used, int a_at, int a_split, int b_at, int b_split, tiling_t *o) {
    if (i == z->na && j == z->nb) return a_split && b_split;
    const int a = z->a[(z->s + i) % z->na], b = z->b[((z->t - j) % z->nb + z->nb) % z->nb];
    if (i < z->na) {
        const int a2 = z->a[(z->s + i + 1) % z->na];
        const int cost = (i + 1 == z->na && j == z->nb) ? 0 : in_face(a2, b);
        if (used + cost <= z->budget && !in_face3(a, a2, b)) {
            push(o, a, a2, b);
            if (zip(z, i + 1, j, used + cost, a_at < 0 ? j : a_at, a_split || (a_at >= 0 && a_at != j), b_at, b_split, o))
                return 1;
            --o->n;
        }
    }
    if (j < z->nb) {
        const int b2 = z->b[((z->t - j - 1) % z->nb + z->nb) % z->nb];
        const int cost = (i == z->na && j + 1 == z->nb) ? 0 : in_face(a, b2);
        if (used + cost <= z->budget && !in_face3(b2, b, a)) {
            push(o, b2, b, a);
            if (zip(z, i, j + 1, used + cost, a_at, a_split, b_at < 0 ? i : b_at, b_split || (b_at >= 0 && b_at != i), o))
                return 1;
            --o->n;
        }
    }
    return 0;
}

/* Tube without interior vertex: na + nb triangles, A walked forwards and B backwards from the bridge A[s] - B[t]. */
static int tube(const int *a, int na, const int *b, int nb, tiling_t *o) {
    zip_t z = {a, b, na, nb, 0, 0, 0};
    for (z.budget = 0; z.budget <= na + nb; ++z.budget)
        for (z.s = 0; z.s < na; ++z.s)
            for (z.t = 0; z.t < nb; ++z.t) {
                const int c0 = in_face(a[z.s], b[z.t]);
                if (c0 > z.budget) continue;
                const int mark = o->n;
                if (zip(&z, 0, 0, c0, -1, 0, -1, 0, o)) return 1;
                o->n = mark;
            }
    return 0;
}

/* Tube through the centre vertex (6.1.2): segments A[k] -> A[k+1] and B[j] -> B[j+1] plus the bridges A[k+1] - B[j] and
 * A[k] - B[j+1] bound a quad (two triangles); the polygon of everything else is fanned around the centre. */
static int tube_centre(const int *a, int na, const int *b, int nb, tiling_t *o) {
    for (int k = 0; k < na; ++k)
        for (int j = 0; j < nb; ++j) {
            const int a0 = a[k], a1 = a[(k + 1) % na], b0 = b[j], b1 = b[(j + 1) % nb];
            if (in_face(a1, b0) || in_face(a0, b1)) continue;
            if (!in_face(a0, b0)) {
                push(o, a0, a1, b0);
                push(o, a0, b0, b1);
            } else if (!in_face(a1, b1)) {
                push(o, a0, a1, b1);
                push(o, a1, b0, b1);
            } else {
                continue;
            }
            int ring[24], n = 0;
            for (int i = 0; i < na; ++i) ring[n++] = a[(k + 1 + i) % na];
            for (int i = 0; i < nb; ++i) ring[n++] = b[(j + 1 + i) % nb];
            fan(ring, n, o);
            return 1;
        }
    return 0;
}

static int label_is(const char *label, const char *const *set) {
    for (; *set; ++set)
        if (!strcmp(label, *set)) return 1;
    return 0;
}

/* The tiling of one (configuration, face resolution, tunnel) triple; returns 0 if it cannot be built. */
static int build_tiling(int config, int pmask, int tunnel, tiling_t *o, char *label_out) {
    static const char *const centre_fan[] = {"7.3", "10.2", "12.2", "13.3", "13.3_", "13.4", NULL};
    geometry_init();
    o->n = 0;
    subcase_t s = subcase(config, pmask);
    if (tunnel && !s.interior) return 0;
    char label[16];
    strcpy(label, s.label);
    if (s.interior) strcat(label, tunnel ? ".2" : ".1");
    if (label_out) strcpy(label_out, label);
    loops_t L = contour_loops(config, pmask);
    int done[4] = {0, 0, 0, 0};
    if (tunnel) {
        int ia = 0, ib = 1;
        if (L.n == 3) { /* 13.5.2: the hexagon and the triangle around the lone positive corner */
            ia = ib = -1;
            for (int l = 0; l < 3; ++l) {
                if (L.len[l] == 6) ia = l;
                else {
                    int around_lone = 1;
                    for (int q = 0; q < 3; ++q)
                        if (g_edge[L.e[l][q]][0] != s.lone && g_edge[L.e[l][q]][1] != s.lone) around_lone = 0;
                    if (around_lone) ib = l;
                }
            }
            if (ia < 0 || ib < 0) return 0;
        } else if (L.n != 2) {
            return 0;
        }
        const int ok = !strcmp(label, "6.1.2") ? tube_centre(L.e[ia], L.len[ia], L.e[ib], L.len[ib], o)
                                               : tube(L.e[ia], L.len[ia], L.e[ib], L.len[ib], o);
        if (!ok) return 0;
        done[ia] = done[ib] = 1;
    }
    int big = -1;
    for (int l = 0; l < L.n; ++l)
        if (!done[l] && (big < 0 || L.len[l] > L.len[big])) big = l;
    for (int l = 0; l < L.n; ++l) {
        if (done[l]) continue;
        if (l == big && label_is(label, centre_fan)) fan(L.e[l], L.len[l], o);
        else if (!disk(L.e[l], 0, L.len[l] - 1, o)) {
            if (strcmp(label, "13.x")) return 0;
            fan(L.e[l], L.len[l], o); /* only for the face-test combinations that cannot occur */
        }
    }
    return 1;
}

/* cache: [config][pmask][tunnel] */
static tiling_t *g_cache[256][64][2];

static const tiling_t *cached_tiling(int config, int pmask, int tunnel) {
    tiling_t **slot = &g_cache[config][pmask][tunnel];
    if (!*slot) {
        tiling_t *t = (tiling_t *)malloc(sizeof(tiling_t));
        if (!t || !build_tiling(config, pmask, tunnel, t, NULL)) {
            fprintf(stderr, "mc_oracle: no tiling for configuration %d, faces %d, tunnel %d\n", config, pmask, tunnel);
            abort();
        }
        *slot = t;
    }
    return *slot;
}

/* Full decision for one cube: face tests, sub-case, interior test. */
static const tiling_t *cube_tiling(int config, const double v[8], int *pmask_out, int *tunnel_out) {
    geometry_init();
    int pmask = 0;
    for (int f = 0; f < NF; ++f)
        if (face_ambiguous(config, f) && face_positives_joined(config, f, v)) pmask |= 1 << f;
    int tunnel = 0;
    static int8_t has_interior[256][64]; /* 0 unknown, 1 no, 2 yes */
    static subcase_t interior_desc[256][64];
    if (!has_interior[config][pmask]) {
        interior_desc[config][pmask] = subcase(config, pmask);
        has_interior[config][pmask] = interior_desc[config][pmask].interior ? 2 : 1;
    }
    if (has_interior[config][pmask] == 2) tunnel = interior_says_tunnel(&interior_desc[config][pmask], v);
    if (pmask_out) *pmask_out = pmask;
    if (tunnel_out) *tunnel_out = tunnel;
    return cached_tiling(config, pmask, tunnel);
}

/* ---- introspection for the tests ----------------------------------------------------------------------------------------- */
/* Corners of oracle face f (cyclic), so that a test can translate its own face numbering. */
void hvo_mc_face_corners(int f, int out[4]) {
    geometry_init();
    for (int k = 0; k < 4; ++k) out[k] = g_face[f][k];
}

/* Tiling of a (configuration, pmask, tunnel) triple: returns the number of triangles (-1: not a valid triple), fills
 * tris[3 * n] with edge ids (12 = centre vertex) and label (>= 16 chars). */
int hvo_mc_tiling(int config, int pmask, int tunnel, uint8_t *tris, char *label) {
    tiling_t t;
    if (!build_tiling(config, pmask, tunnel, &t, label)) return -1;
    memcpy(tris, t.t, (size_t)t.n * 3);
    return t.n;
}

/* Decision for one cube given its eight corner values (already minus the level), Lewiner corner order. */
int hvo_mc_cube(const double v[8], int *config_out, int *pmask_out, int *tunnel_out, uint8_t *tris, char *label) {
    int config = 0;
    for (int i = 0; i < NC; ++i)
        if (v[i] > 0.0) config |= 1 << i;
    int pmask = 0, tunnel = 0;
    geometry_init();
    if (config != 0 && config != 255) cube_tiling(config, v, &pmask, &tunnel);
    *config_out = config, *pmask_out = pmask, *tunnel_out = tunnel;
    return hvo_mc_tiling(config, pmask, tunnel, tris, label);
}

/* ---- the extraction -------------------------------------------------------------------------------------------------------- */
typedef struct {
    float *verts;      /* nv x 3, voxel units, (axis0, axis1, axis2) */
    float *normals;    /* nv x 3 */
    int64_t *keys;     /* nv: (linear index of the owning voxel) * 4 + axis (3 = centre vertex of its cube) */
    int32_t *faces;    /* nf x 3 */
    int64_t *cubes;    /* nf: linear index of the cube each face came from */
    int64_t nv, nf, cap_v, cap_f;
} hvo_mesh;

static int grow(void **p, int64_t *cap, int64_t need, size_t elem) {
    if (need <= *cap) return 0;
    int64_t c = *cap ? *cap : 1024;
    while (c < need) c *= 2;
    void *q = realloc(*p, (size_t)c * elem);
    if (!q) return -1;
    *p = q;
    *cap = c;
    return 0;
}

void hvo_mc_free(hvo_mesh *m) {
    if (!m) return;
    free(m->verts);
    free(m->normals);
    free(m->keys);
    free(m->faces);
    free(m->cubes);
    free(m);
}

/* -1 = level outside [min, max] (skimage raises ValueError), -2 = out of memory, -3 = bad dims;
 * otherwise returns the mesh (possibly with nv == 0: skimage raises RuntimeError then). */
int hvo_marching_cubes(const float *vol, int64_t X, int64_t Y, int64_t Z, double level, hvo_mesh **out) {
    *out = NULL;
    if (X < 2 || Y < 2 || Z < 2) return -3;
    geometry_init();
    const int64_t n = X * Y * Z;
    float mn = vol[0], mx = vol[0];
    for (int64_t i = 1; i < n; ++i) {
        if (vol[i] < mn) mn = vol[i];
        if (vol[i] > mx) mx = vol[i];
    }
    if (level < (double)mn || level > (double)mx) return -1;

    hvo_mesh *m = (hvo_mesh *)calloc(1, sizeof(hvo_mesh));
    if (!m) return -2;
    int64_t cap_n = 0, cap_k = 0, cap_c = 0;
    /* two sliding layers (owner x == current x, current x + 1) of [Y][Z][4] vertex indices */
    const int64_t layer_n = Y * Z * 4; /* 3 edges + the centre vertex of the cube based at the voxel */
    int32_t *layer[2];
    layer[0] = (int32_t *)malloc((size_t)layer_n * sizeof(int32_t));
    layer[1] = (int32_t *)malloc((size_t)layer_n * sizeof(int32_t));
    if (!layer[0] || !layer[1]) return -2;
    memset(layer[0], 0xff, (size_t)layer_n * sizeof(int32_t));
    memset(layer[1], 0xff, (size_t)layer_n * sizeof(int32_t));

    int corner_at[2][2][2];
    for (int i = 0; i < 8; ++i) corner_at[g_corner[i][0]][g_corner[i][1]][g_corner[i][2]] = i;

    for (int64_t x = 0; x + 1 < X; ++x) {
        if (x > 0) { /* slide: the upper layer becomes the lower one */
            int32_t *t = layer[0];
            layer[0] = layer[1];
            layer[1] = t;
            memset(layer[1], 0xff, (size_t)layer_n * sizeof(int32_t));
        }
        for (int64_t y = 0; y + 1 < Y; ++y) {
            for (int64_t z = 0; z + 1 < Z; ++z) {
                double v[8];
                int index = 0;
                for (int i = 0; i < 8; ++i) {
                    const int64_t cx = x + g_corner[i][0], cy = y + g_corner[i][1], cz = z + g_corner[i][2];
                    v[i] = (double)vol[(cx * Y + cy) * Z + cz] - level;
                    if (v[i] > 0.0) index |= 1 << i;
                }
                if (index == 0 || index == 255) continue;
                const tiling_t *tl = cube_tiling(index, v, NULL, NULL);
                const int ntri = tl->n;
                const uint8_t *tri = &tl->t[0][0];
                if (grow((void **)&m->faces, &m->cap_f, (m->nf + ntri) * 3, sizeof(int32_t)) ||
                    grow((void **)&m->cubes, &cap_c, m->nf + ntri, sizeof(int64_t)))
                    return -2;
                for (int t = 0; t < ntri; ++t) m->cubes[m->nf + t] = (x * Y + y) * Z + z;
                /* weights and down-hill in-cube gradients of the 8 corners */
                double w[8], g[8][3];
                for (int i = 0; i < 8; ++i) {
                    w[i] = 1.0 / ((double)FLT_EPSILON + fabs(v[i]));
                    for (int k = 0; k < 3; ++k) {
                        int o0[3] = {g_corner[i][0], g_corner[i][1], g_corner[i][2]};
                        int o1[3] = {o0[0], o0[1], o0[2]};
                        o0[k] = 0;
                        o1[k] = 1;
                        g[i][k] = v[corner_at[o0[0]][o0[1]][o0[2]]] - v[corner_at[o1[0]][o1[1]][o1[2]]];
                    }
                }
                for (int t = 0; t < 3 * ntri; ++t) {
                    const int e = tri[t];
                    int axis = 3, own[3] = {0, 0, 0}, lo = 0, hi = 0; /* axis 3: centre vertex, owned by the cube's base voxel */
                    if (e < 12) {
                        lo = g_edge[e][0], hi = g_edge[e][1];
                        axis = g_edge_axis[e];
                        if (g_corner[lo][axis] > g_corner[hi][axis]) { const int s = lo; lo = hi; hi = s; }
                        for (int k = 0; k < 3; ++k) own[k] = g_corner[lo][k];
                    }
                    int32_t *slot = layer[own[0]] + ((y + own[1]) * Z + (z + own[2])) * 4 + axis;
                    if (*slot < 0) {
                        if (grow((void **)&m->verts, &m->cap_v, (m->nv + 1) * 3, sizeof(float)) ||
                            grow((void **)&m->normals, &cap_n, (m->nv + 1) * 3, sizeof(float)) ||
                            grow((void **)&m->keys, &cap_k, m->nv + 1, sizeof(int64_t)))
                            return -2;
                        const int64_t base[3] = {x + own[0], y + own[1], z + own[2]};
                        float *p = m->verts + 3 * m->nv;
                        if (e < 12) {
                            for (int k = 0; k < 3; ++k) p[k] = (float)base[k];
                            p[axis] = (float)((double)base[axis] + w[hi] / (w[lo] + w[hi]));
                        } else { /* centre of mass of the 8 corners */
                            double f[3] = {0.0, 0.0, 0.0}, ff = 0.0;
                            for (int i = 0; i < 8; ++i) {
                                for (int k = 0; k < 3; ++k) f[k] += (double)g_corner[i][k] * w[i];
                                ff += w[i];
                            }
                            for (int k = 0; k < 3; ++k) p[k] = (float)((double)base[k] + f[k] / ff);
                        }
                        m->normals[3 * m->nv] = m->normals[3 * m->nv + 1] = m->normals[3 * m->nv + 2] = 0.0f;
                        m->keys[m->nv] = ((base[0] * Y + base[1]) * Z + base[2]) * 4 + axis;
                        *slot = (int32_t)m->nv++;
                    }
                    m->faces[3 * m->nf + t] = *slot;
                    /* this cube's gradient at the vertex, added once per triangle corner that references it */
                    float *nrm = m->normals + 3 * (size_t)*slot;
                    for (int k = 0; k < 3; ++k) {
                        double gk;
                        if (e < 12) {
                            gk = g[lo][k] * w[lo] + g[hi][k] * w[hi];
                        } else {
                            gk = 0.0;
                            for (int i = 0; i < 8; ++i) gk += w[i] * g[i][k];
                        }
                        nrm[k] += (float)gk;
                    }
                }
                m->nf += ntri;
            }
        }
    }
    free(layer[0]);
    free(layer[1]);
    for (int64_t i = 0; i < m->nv; ++i) {
        float *p = m->normals + 3 * i;
        const double len = sqrt((double)p[0] * p[0] + (double)p[1] * p[1] + (double)p[2] * p[2]);
        if (len > 0.0)
            for (int k = 0; k < 3; ++k) p[k] = (float)((double)p[k] / len);
    }
    *out = m;
    return 0;
}

int64_t hvo_mesh_nv(const hvo_mesh *m) { return m->nv; }
int64_t hvo_mesh_nf(const hvo_mesh *m) { return m->nf; }
void hvo_mesh_copy(const hvo_mesh *m, float *verts, float *normals, int64_t *keys, int32_t *faces, int64_t *cubes) {
    if (cubes) memcpy(cubes, m->cubes, (size_t)m->nf * sizeof(int64_t));
    memcpy(verts, m->verts, (size_t)m->nv * 3 * sizeof(float));
    memcpy(normals, m->normals, (size_t)m->nv * 3 * sizeof(float));
    memcpy(keys, m->keys, (size_t)m->nv * sizeof(int64_t));
    memcpy(faces, m->faces, (size_t)m->nf * 3 * sizeof(int32_t));
}
