/*
 * C restatement of the reference's site-state assignment -- TEST INFRASTRUCTURE ONLY.
 *
 * Follows reference src/gemdat/transitions.py:736-809 (and the MDAnalysis 2.6.1 /
 * scipy cKDTree arithmetic it reaches, restated in oracle/mda.py) as a brute-force
 * loop over every (atom-frame, site) pair: float64 frac*L -> float32 -> wrap
 * (_ortho_pbc) -> optional periodic image per axis (+L if x <= cutoff, else -L if
 * L - x <= cutoff) -> float64 dx*dx + dy*dy + dz*dz <= r*r (x, y, z order, every
 * operation rounded; compile with -ffp-contract=off).  Last match in overwrite
 * order wins.  Used by the tests to check large inputs quickly and by bench.py as
 * an optional multi-threaded CPU baseline ("port").  Orthorhombic, diagonal
 * lattices only.
 */
#include <math.h>
#include <stdint.h>
#include <pthread.h>
#include <stdlib.h>

static float wrap_ortho(float c, float L) {
    const double inv = 1.0 / (double)L;
    if (c < 0.0f) {
        float once = c + L;
        if (once < 0.0f) {
            int s = (int)floor((double)c * inv);
            c = c - (float)s * L;
            if (c < 0.0f) c = c + L;
        } else {
            c = once;
        }
    }
    if (c >= L) {
        float once = c - L;
        if (once >= L) {
            int s = (int)floor((double)c * inv);
            c = c - (float)s * L;
            if (c >= L) c = c - L;
        } else {
            c = once;
        }
    }
    return c;
}

/*
 * pos        [n,3] fractional float64
 * diag       [3]   lattice lengths (diagonal matrix)
 * site32     [n_ranked,3] staged (float32, wrapped) site centres in overwrite order
 * natural    [n_ranked]   value written for a hit (already remapped by the caller if needed)
 * r2_outer/r2_inner [n_ranked]
 * out_outer/out_inner [n] int64, -1 = no site
 * visited    [2*n_ranked] uint8 (may be NULL)
 * n_threads  POSIX threads over contiguous atom-frame ranges
 */
typedef struct {
    const double *pos;
    int64_t begin, end;
    const double *diag;
    double cutoff;
    const float *site32;
    const int64_t *natural;
    const double *r2_outer, *r2_inner;
    int64_t n_ranked;
    int64_t *out_outer, *out_inner;
    uint8_t *visited;
} job_t;

static void *run_job(void *arg) {
    const job_t *jb = (const job_t *)arg;
    float box[3];
    for (int j = 0; j < 3; ++j) box[j] = (float)jb->diag[j];
    const float r32 = (float)jb->cutoff;
    const int64_t n_ranked = jb->n_ranked;
    for (int64_t i = jb->begin; i < jb->end; ++i) {
        double a0[3], a1[3];
        int has[3];
        for (int j = 0; j < 3; ++j) {
            const double cart = jb->pos[3 * i + j] * jb->diag[j];
            const float c = wrap_ortho((float)cart, box[j]);
            a0[j] = (double)c;
            has[j] = 1;
            if (c <= r32) {
                a1[j] = (double)(float)(c + box[j]);
            } else if ((float)(box[j] - c) <= r32) {
                a1[j] = (double)(float)(c - box[j]);
            } else {
                has[j] = 0;
                a1[j] = 0.0;
            }
        }
        int64_t bo = -1, bi = -1;
        for (int64_t e = 0; e < n_ranked; ++e) {
            double q[3];
            for (int j = 0; j < 3; ++j) {
                const double s = (double)jb->site32[3 * e + j];
                const double d = s - a0[j];
                q[j] = d * d;
                if (has[j]) {
                    const double d1 = s - a1[j];
                    const double q1 = d1 * d1;
                    if (q1 < q[j]) q[j] = q1;
                }
            }
            double d2 = q[0] + q[1];
            d2 = d2 + q[2];
            if (d2 <= jb->r2_outer[e]) {
                bo = jb->natural[e];
                if (jb->visited) jb->visited[e] = 1;
            }
            if (d2 <= jb->r2_inner[e]) {
                bi = jb->natural[e];
                if (jb->visited) jb->visited[n_ranked + e] = 1;
            }
        }
        jb->out_outer[i] = bo;
        jb->out_inner[i] = bi;
    }
    return NULL;
}

void oracle_states_bf(const double *pos, int64_t n, const double *diag, double cutoff,
                      const float *site32, const int64_t *natural, const double *r2_outer,
                      const double *r2_inner, int64_t n_ranked, int64_t *out_outer,
                      int64_t *out_inner, uint8_t *visited, int n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t tid[256];
    job_t jobs[256];
    for (int t = 0; t < n_threads; ++t) {
        job_t jb = {pos, n * t / n_threads, n * (t + 1) / n_threads, diag, cutoff, site32, natural,
                    r2_outer, r2_inner, n_ranked, out_outer, out_inner, visited};
        jobs[t] = jb;
        pthread_create(&tid[t], NULL, run_job, &jobs[t]);
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(tid[t], NULL);
}
