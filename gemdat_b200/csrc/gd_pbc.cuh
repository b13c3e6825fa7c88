// Minimum-image geometry shared by gd_pbc_distances and gd_rdf_hist.
//
// Follows pymatgen's Lattice.get_all_distances -> pbc_shortest_vectors (coord_cython.pyx,
// pymatgen 2024.1.27, restated from memory -- the source is not vendored by the reference):
//   1. both coordinate sets are mapped into the LLL-REDUCED basis (Lattice.lll_matrix /
//      lll_inverse, delta = 0.75) and wrapped to [0,1) there; in a skewed or non-reduced cell
//      the 27 neighbouring images of the RAW basis do not contain the minimum image;
//   2. cart = frac_lll @ lll_matrix per point, pre = cart2 - cart1;
//   3. d2 = min over the 27 images of |pre + image @ lll_matrix|^2, summed (x^2 + y^2) + z^2.
// The reduction is done on the host (3x3), the kernels get the two matrices and the 27
// Cartesian image vectors.
#pragma once

#include <math.h>
#include <string.h>

struct GdPbc {
    double inv[9];      // lll_inverse, row-major: frac_lll = frac @ inv
    double m[9];        // lll_matrix, rows = reduced basis vectors (Angstrom)
    double img[27][3];  // image vectors (i, j, k) in (-1, 0, 1)^3, k fastest, @ lll_matrix
    double axis_len[3]; // axis_aligned: length of the basis vector along Cartesian axis c
    int axis_aligned;   // every reduced basis vector lies along one Cartesian axis
};

namespace gd_pbc_detail {

inline void gram_schmidt(const double b[3][3], double bs[3][3], double mu[3][3], double norm[3]) {
    for (int i = 0; i < 3; ++i) {
        for (int c = 0; c < 3; ++c) bs[i][c] = b[i][c];
        for (int j = 0; j < i; ++j) {
            double dot = 0.0;
            for (int c = 0; c < 3; ++c) dot += b[i][c] * bs[j][c];
            mu[i][j] = norm[j] > 0.0 ? dot / norm[j] : 0.0;
            for (int c = 0; c < 3; ++c) bs[i][c] -= mu[i][j] * bs[j][c];
        }
        norm[i] = 0.0;
        for (int c = 0; c < 3; ++c) norm[i] += bs[i][c] * bs[i][c];
    }
}

}  // namespace gd_pbc_detail

// LLL reduction of the three lattice rows (delta = 0.75, nearest-integer size reduction with
// ties to even like Python's round).  lll = map @ lattice with `map` unimodular.
inline void gd_lll_reduce(const double *lattice, double lll[9], double map[9]) {
    using namespace gd_pbc_detail;
    double b[3][3], mp[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    memcpy(b, lattice, sizeof(b));
    double bs[3][3], mu[3][3] = {{0}}, norm[3];
    const double delta = 0.75;
    int k = 1;
    for (int guard = 0; k < 3 && guard < 1000; ++guard) {
        gram_schmidt(b, bs, mu, norm);
        for (int j = k - 1; j >= 0; --j) {
            const double q = nearbyint(mu[k][j]);
            if (q != 0.0) {
                for (int c = 0; c < 3; ++c) {
                    b[k][c] -= q * b[j][c];
                    mp[k][c] -= q * mp[j][c];
                }
                gram_schmidt(b, bs, mu, norm);
            }
        }
        if (norm[k] >= (delta - mu[k][k - 1] * mu[k][k - 1]) * norm[k - 1]) {
            ++k;
        } else {
            for (int c = 0; c < 3; ++c) {
                double t = b[k][c];
                b[k][c] = b[k - 1][c];
                b[k - 1][c] = t;
                t = mp[k][c];
                mp[k][c] = mp[k - 1][c];
                mp[k - 1][c] = t;
            }
            k = k > 1 ? k - 1 : 1;
        }
    }
    memcpy(lll, b, sizeof(b));
    memcpy(map, mp, sizeof(mp));
}

inline GdPbc gd_pbc_build(const double *h_lattice) {
    GdPbc g;
    double map[9];
    gd_lll_reduce(h_lattice, g.m, map);
    // inverse of the unimodular integer mapping: adjugate / det, exact
    const double *a = map;
    const double det = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) +
                       a[2] * (a[3] * a[7] - a[4] * a[6]);
    const double s = det < 0 ? -1.0 : 1.0;
    g.inv[0] = s * (a[4] * a[8] - a[5] * a[7]);
    g.inv[1] = s * (a[2] * a[7] - a[1] * a[8]);
    g.inv[2] = s * (a[1] * a[5] - a[2] * a[4]);
    g.inv[3] = s * (a[5] * a[6] - a[3] * a[8]);
    g.inv[4] = s * (a[0] * a[8] - a[2] * a[6]);
    g.inv[5] = s * (a[2] * a[3] - a[0] * a[5]);
    g.inv[6] = s * (a[3] * a[7] - a[4] * a[6]);
    g.inv[7] = s * (a[1] * a[6] - a[0] * a[7]);
    g.inv[8] = s * (a[0] * a[4] - a[1] * a[3]);
    for (int k = 0; k < 9; ++k) g.inv[k] = nearbyint(g.inv[k]) + 0.0;  // (+0.0: no negative zero)
    int n = 0;
    for (int i = -1; i <= 1; ++i)
        for (int j = -1; j <= 1; ++j)
            for (int k = -1; k <= 1; ++k, ++n)
                for (int c = 0; c < 3; ++c)  // dot_2d: ((0 + i m0c) + j m1c) + k m2c
                    g.img[n][c] = ((0.0 + i * g.m[c]) + j * g.m[3 + c]) + k * g.m[6 + c];
    // axis aligned: each row has one non-zero entry and the rows use different axes
    g.axis_aligned = 1;
    int used[3] = {0, 0, 0};
    for (int c = 0; c < 3; ++c) g.axis_len[c] = 0.0;
    for (int r = 0; r < 3; ++r) {
        int nz = 0, ax = -1;
        for (int c = 0; c < 3; ++c)
            if (g.m[3 * r + c] != 0.0) {
                ++nz;
                ax = c;
            }
        if (nz != 1 || used[ax]) {
            g.axis_aligned = 0;
            break;
        }
        used[ax] = 1;
        g.axis_len[ax] = fabs(g.m[3 * r + ax]);  // (the image set is symmetric: the sign is irrelevant)
    }
    return g;
}

#ifdef __CUDACC__

// Cartesian coordinates (reduced basis) of a fractional point wrapped to [0,1) there
__device__ __forceinline__ void gd_pbc_cart(const GdPbc &g, double f0, double f1, double f2,
                                            double out[3]) {
    double w[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double v = (f0 * g.inv[k] + f1 * g.inv[3 + k]) + f2 * g.inv[6 + k];
        w[k] = v - floor(v);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) out[k] = (w[0] * g.m[k] + w[1] * g.m[3 + k]) + w[2] * g.m[6 + k];
}

// squared minimum-image distance of two points given by gd_pbc_cart
__device__ __forceinline__ double gd_pbc_d2(const GdPbc &g, const double a[3], const double b[3]) {
    const double p0 = b[0] - a[0], p1 = b[1] - a[1], p2 = b[2] - a[2];
    if (g.axis_aligned) {
        // the 27 image vectors are (+-L or 0) per Cartesian axis: the minimum separates per
        // axis (rounding is monotonic: min over images of fl(fl(X + Y) + Z) =
        // fl(fl(min X + min Y) + min Z)), and fl(v^2) is monotonic in |v|, so per axis the
        // candidate of smallest magnitude among p - L, p, p + L is squared once
        // Both points are wrapped into the cell, so |p| < L and p + sign(p) L is never the
        // shortest: the candidates are p and q = p - sign(p) L.
        const double pre[3] = {p0, p1, p2};
        double sq[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double q = pre[c] - copysign(g.axis_len[c], pre[c]);
            const double v = fabs(pre[c]) <= fabs(q) ? pre[c] : q;
            sq[c] = v * v;
        }
        return (sq[0] + sq[1]) + sq[2];
    }
    double best = 1e300;
#pragma unroll
    for (int n = 0; n < 27; ++n) {
        const double da = p0 + g.img[n][0], db = p1 + g.img[n][1], dc = p2 + g.img[n][2];
        best = fmin(best, (da * da + db * db) + dc * dc);
    }
    return best;
}

#endif  // __CUDACC__
