/*
 * oracle/nnd_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's brute-force 3-D nearest-neighbour
 * ("nnd" / chamfer) forward and backward.  It exists so that tests/, the
 * smoke() check and bench.py's cpu_baseline / --impl reference leg have
 * something to compare the CUDA path with.  Nothing in the product path
 * (pointcloudregistration_b200/, torch_nndistance/) may import, link or call
 * it; the product path fails loudly when the CUDA library is missing.
 *
 * Parity pinning: the reference ships no golden vectors for this path
 * (dip/torch-nndistance/test.py only prints).  This restatement is pinned
 * by running the reference's own CPU source unmodified (oracle/_ref, built by
 * oracle/build_ref.py from /root/reference/dip/torch-nndistance/src/my_lib.cpp)
 * on seeded inputs and (a) comparing bit for bit in tests/test_oracle.py when
 * oracle/_ref is present, (b) committing SHA-256 digests + small raw vectors of
 * those reference outputs under tests/golden/ (tests/golden/make_golden.py).
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared -pthread  (NEVER -march=native
 * or -ffast-math without -ffp-contract=off: FMA contraction changes dist bits).
 *
 * Semantics followed (reference file:line):
 *   forward  : dip/torch-nndistance/src/my_lib.cpp:3-25   (nnsearch)
 *              dip/torch-nndistance/src/my_lib.cpp:56-57  (called twice, roles swapped)
 *   backward : dip/torch-nndistance/src/my_lib.cpp:91-130
 */
#include <pthread.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

/* One direction, one batch element: for each query j the lowest-index target
 * k minimising fl(fl(fl(dx*dx)+fl(dy*dy))+fl(dz*dz)), d = target - query.
 * my_lib.cpp:5-22.  `best` is kept as float: the reference widens the float
 * value to double for the compare only, which orders identically. */
static void nn_rows(const float *q, size_t nq, const float *t, size_t nt,
                    float *dist, int32_t *idx)
{
    for (size_t j = 0; j < nq; ++j) {
        const float qx = q[3 * j], qy = q[3 * j + 1], qz = q[3 * j + 2];
        float best = 0.0f;              /* my_lib.cpp:9  (m == 0 -> 0, 0) */
        int32_t besti = 0;              /* my_lib.cpp:10 */
        for (size_t k = 0; k < nt; ++k) {
            const float dx = t[3 * k] - qx;          /* my_lib.cpp:12 */
            const float dy = t[3 * k + 1] - qy;      /* my_lib.cpp:13 */
            const float dz = t[3 * k + 2] - qz;      /* my_lib.cpp:14 */
            const float xx = dx * dx, yy = dy * dy, zz = dz * dz;
            const float s = xx + yy;                 /* my_lib.cpp:15: (x*x+y*y)+z*z */
            const float d = s + zz;
            if (k == 0 || d < best) {                /* my_lib.cpp:16 strict <, first wins */
                best = d;
                besti = (int32_t)k;
            }
        }
        dist[j] = best;                 /* my_lib.cpp:21 */
        idx[j] = besti;                 /* my_lib.cpp:22 */
    }
}

/* One-direction query over a batch: q[b,nq,3] against t[b,nt,3]. */
int nnd_oracle_query_f32(const float *q, const float *t, int b, int nq, int nt,
                         float *dist, int32_t *idx)
{
    for (int i = 0; i < b; ++i)
        nn_rows(q + (size_t)i * nq * 3, (size_t)nq, t + (size_t)i * nt * 3, (size_t)nt,
                dist + (size_t)i * nq, idx + (size_t)i * nq);
    return 0;
}

/* my_lib.cpp:28-60 (nnd_forward): direction 1 then direction 2. */
int nnd_oracle_forward_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                           float *dist1, float *dist2, int32_t *idx1, int32_t *idx2)
{
    nnd_oracle_query_f32(xyz1, xyz2, b, n, m, dist1, idx1);
    nnd_oracle_query_f32(xyz2, xyz1, b, m, n, dist2, idx2);
    return 0;
}

/* my_lib.cpp:64-133 (nnd_backward).  Sequential accumulation order:
 * zero-fill, then per batch the dist1 term (ascending j) then the dist2 term. */
int nnd_oracle_backward_f32(const float *xyz1, const float *xyz2,
                            const float *graddist1, const float *graddist2,
                            const int32_t *idx1, const int32_t *idx2,
                            int b, int n, int m, float *gradxyz1, float *gradxyz2)
{
    const size_t N = (size_t)n, M = (size_t)m;
    for (size_t i = 0; i < (size_t)b * N * 3; ++i) gradxyz1[i] = 0.0f;  /* :91-92 */
    for (size_t i = 0; i < (size_t)b * M * 3; ++i) gradxyz2[i] = 0.0f;  /* :93-94 */
    for (size_t i = 0; i < (size_t)b; ++i) {
        const float *p1 = xyz1 + i * N * 3, *p2 = xyz2 + i * M * 3;
        float *g1 = gradxyz1 + i * N * 3, *g2 = gradxyz2 + i * M * 3;
        for (size_t j = 0; j < N; ++j) {                               /* :96-113 */
            const size_t j2 = (size_t)idx1[i * N + j];
            const float g = graddist1[i * N + j] * 2;                  /* :105 */
            for (int c = 0; c < 3; ++c) {
                const float diff = p1[3 * j + c] - p2[3 * j2 + c];
                const float tt = g * diff;
                g1[3 * j + c] += tt;                                   /* :107-109 */
                g2[3 * j2 + c] -= tt;                                  /* :110-112 */
            }
        }
        for (size_t j = 0; j < M; ++j) {                               /* :114-129 */
            const size_t j2 = (size_t)idx2[i * M + j];
            const float g = graddist2[i * M + j] * 2;                  /* :122 */
            for (int c = 0; c < 3; ++c) {
                const float diff = p2[3 * j + c] - p1[3 * j2 + c];
                const float tt = g * diff;
                g2[3 * j + c] += tt;                                   /* :123-125 */
                g1[3 * j2 + c] -= tt;                                  /* :126-128 */
            }
        }
    }
    return 0;
}

/* ---- all-core driver: the same nn_rows loop over disjoint query slices ---- */
typedef struct {
    const float *q, *t;
    size_t nq, nt;
    float *dist;
    int32_t *idx;
} slice_t;

static void *slice_main(void *p)
{
    slice_t *s = (slice_t *)p;
    nn_rows(s->q, s->nq, s->t, s->nt, s->dist, s->idx);
    return NULL;
}

/* Same result as nnd_oracle_query_f32 (rows are independent); `threads`
 * workers each take a contiguous slice of every batch element's queries. */
int nnd_oracle_query_mt_f32(const float *q, const float *t, int b, int nq, int nt,
                            float *dist, int32_t *idx, int threads)
{
    if (threads < 1) threads = 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    slice_t *sl = (slice_t *)malloc(sizeof(slice_t) * (size_t)threads);
    if (!th || !sl) { free(th); free(sl); return 1; }
    for (int i = 0; i < b; ++i) {
        const size_t per = ((size_t)nq + (size_t)threads - 1) / (size_t)threads;
        int started = 0;
        for (int w = 0; w < threads; ++w) {
            const size_t lo = per * (size_t)w;
            if (lo >= (size_t)nq) break;
            const size_t hi = lo + per < (size_t)nq ? lo + per : (size_t)nq;
            sl[w].q = q + ((size_t)i * nq + lo) * 3;
            sl[w].nq = hi - lo;
            sl[w].t = t + (size_t)i * nt * 3;
            sl[w].nt = (size_t)nt;
            sl[w].dist = dist + (size_t)i * nq + lo;
            sl[w].idx = idx + (size_t)i * nq + lo;
            if (pthread_create(&th[w], NULL, slice_main, &sl[w]) != 0) { slice_main(&sl[w]); th[w] = 0; }
            started = w + 1;
        }
        for (int w = 0; w < started; ++w)
            if (th[w]) pthread_join(th[w], NULL);
    }
    free(th);
    free(sl);
    return 0;
}

int nnd_oracle_forward_mt_f32(const float *xyz1, const float *xyz2, int b, int n, int m,
                              float *dist1, float *dist2, int32_t *idx1, int32_t *idx2,
                              int threads)
{
    nnd_oracle_query_mt_f32(xyz1, xyz2, b, n, m, dist1, idx1, threads);
    nnd_oracle_query_mt_f32(xyz2, xyz1, b, m, n, dist2, idx2, threads);
    return 0;
}

const char *nnd_oracle_version(void) { return "nnd-oracle 1 (restates my_lib.cpp:3-25,91-130)"; }
