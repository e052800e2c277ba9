/* ORACLE -- test infrastructure, not product code.
 *
 * Plain-C restatement of the reference's CPU TSDF integration
 * (third_party/tsdf_fusion_python/fusion.py:260-303 with the numba helpers :178-215 and rigid_transform :353-358),
 * one voxel at a time, OpenMP over x.  It follows SURVEY.md App. A item by item (same list as
 * oracle/fusion_oracle.py, against which -- and against the reference itself -- tests/ pin it).
 *
 * Used by tests/ as the fast checker at sizes numpy cannot finish in seconds, and by bench.py's `cpu_baseline` /
 * `--impl reference` legs as the timed CPU port of the reference path ("kind": "port").  Never linked into the
 * product library.
 *
 * Build: see oracle/Makefile (gcc -O2 -fopenmp -ffp-contract=off: no FMA contraction, so every mul/add rounds
 * exactly where the reference's numpy / numba code rounds).
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int hvo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to every rank; the timed CPU arm asks for all host threads explicitly. */
void hvo_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* fusion.py:293-295 -- float32 unpack of the b*65536+g*256+r colour */
static inline void unpack(float c, float *r, float *g, float *b) {
    *b = floorf(c / 65536.0f);
    *g = floorf((c - *b * 65536.0f) / 256.0f);
    *r = c - *b * 65536.0f - *g * 256.0f;
}

/*
 * tsdf/weight/color: dense C-order [X][Y][Z] float32 planes (fusion.py:54-57).
 * wx/wy/wz: per-axis float32 world coordinates, A.2 (fusion.py:183-188).
 * w2c: inv(cam_pose) rows 0..2 (3x4 float64, row-major), A.3.  intr: fx, fy, cx, cy as float32, A.4.
 * depth: (H,W) float64 -- the CPU path never narrows depth_im (fusion.py:276-281; demo.py:46 feeds float64, HIVE feeds
 * float32, hive/io.py:1032-1039: widening those is exact); color: (H,W) float32 already folded (A.1, fusion.py:231-232).
 * Returns the number of voxels updated (valid_pts.sum(), fusion.py:279).
 */
long long hvo_integrate(float *tsdf, float *weight, float *color,
                        long long X, long long Y, long long Z,
                        const float *wx, const float *wy, const float *wz,
                        const double *w2c, const float *intr,
                        const double *depth, const float *folded, int H, int W,
                        double trunc, double obs_weight) {
    const double fx = (double)intr[0], fy = (double)intr[1], cx = (double)intr[2], cy = (double)intr[3];
    const float ow32 = (float)obs_weight;
    long long updated = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : updated)
    for (long long x = 0; x < X; ++x) {
        const double px = (double)wx[x];
        for (long long y = 0; y < Y; ++y) {
            const double py = (double)wy[y];
            float *tc = tsdf + (x * Y + y) * Z, *wc = weight + (x * Y + y) * Z, *cc = color + (x * Y + y) * Z;
            for (long long z = 0; z < Z; ++z) {
                const double pz = (double)wz[z];
                /* A.3: left-to-right float64 sums */
                const double qx = w2c[0] * px + w2c[1] * py + w2c[2] * pz + w2c[3];
                const double qy = w2c[4] * px + w2c[5] * py + w2c[6] * pz + w2c[7];
                const double qz = w2c[8] * px + w2c[9] * py + w2c[10] * pz + w2c[11];
                /* A.4: multiply, divide, add; round half to even (rint under the default rounding mode) */
                const double u = rint((qx * fx) / qz + cx);
                const double v = rint((qy * fy) / qz + cy);
                /* A.5 */
                if (!(u >= 0.0 && u < (double)W && v >= 0.0 && v < (double)H && qz > 0.0)) continue;
                const size_t pix = (size_t)((long long)v * W + (long long)u);
                const double d = depth[pix];
                const double diff = d - qz;
                if (!(d > 0.0 && diff >= -trunc)) continue;
                double dist = diff / trunc;
                if (dist > 1.0) dist = 1.0;
                /* A.6 */
                const float w_old = wc[z];
                const float w_new = (float)((double)w_old + obs_weight);
                const float prod = w_old * tc[z];
                tc[z] = (float)(((double)prod + obs_weight * dist) / (double)w_new);
                wc[z] = w_new;
                /* A.7: all float32, half-even */
                float o_r, o_g, o_b, n_r, n_g, n_b;
                unpack(cc[z], &o_r, &o_g, &o_b);
                unpack(folded[pix], &n_r, &n_g, &n_b);
                float b = rintf((w_old * o_b + ow32 * n_b) / w_new);
                float g = rintf((w_old * o_g + ow32 * n_g) / w_new);
                float r = rintf((w_old * o_r + ow32 * n_r) / w_new);
                if (b > 255.0f) b = 255.0f;
                if (g > 255.0f) g = 255.0f;
                if (r > 255.0f) r = 255.0f;
                cc[z] = b * 65536.0f + g * 256.0f + r;
                ++updated;
            }
        }
    }
    return updated;
}
