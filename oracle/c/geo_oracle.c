/* geo_oracle.c -- plain-C restatement of the published algorithms behind the reference's hot path.
 * TEST INFRASTRUCTURE ONLY (checker for tests/, smoke() and bench.py's cpu_baseline); never linked
 * into or called from the product.
 *
 * The reference (m4rch1n0/vqvae) delegates its arithmetic to libraries that are not vendored:
 *   scikit-learn 1.9.0  NearestNeighbors brute force  (src/geo/knn_graph_optimized.py:39-42)
 *   SciPy 1.18.1        csgraph.dijkstra              (src/geo/geo_shortest_paths.py:36-49)
 *   NumPy 2.3.5         float32 pairwise sum, RandomState.choice (src/geo/kmeans_optimized.py:56-61)
 * Each function below restates that library's documented algorithm; tests/test_oracle_c.py pins them
 * against the golden vectors generated from the unmodified reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- exact kNN: squared Euclidean distance in fp64, ties -> lower index (sklearn's heap keeps the
 * earlier index on equal distance).  out_idx[M*kq] ascending by (d2, idx), self included. ---- */
void oracle_knn_bruteforce(const float* z, int64_t M, int d, int kq, int32_t* out_idx, double* out_d2) {
    for (int64_t q = 0; q < M; ++q) {
        int32_t* bi = out_idx + q * kq;
        double* bd = out_d2 + q * kq;
        int cnt = 0;
        for (int64_t c = 0; c < M; ++c) {
            double s = 0.0;
            for (int i = 0; i < d; ++i) {
                double t = (double)z[q * d + i] - (double)z[c * d + i];
                s = fma(t, t, s);
            }
            if (cnt == kq && !(s < bd[kq - 1])) continue; /* reject val >= current worst */
            int pos = cnt < kq ? cnt : kq - 1;
            while (pos > 0 && bd[pos - 1] > s) { /* stable: equal keys keep the earlier index first */
                bd[pos] = bd[pos - 1];
                bi[pos] = bi[pos - 1];
                --pos;
            }
            bd[pos] = s;
            bi[pos] = (int32_t)c;
            if (cnt < kq) ++cnt;
        }
    }
}

/* ---- Dijkstra with a binary heap, fp64 path sums accumulated source -> target exactly like
 * scipy.sparse.csgraph.dijkstra (weights up-cast to DTYPE=float64).  dist[n] = +inf when unreachable. ---- */
typedef struct {
    double key;
    int32_t v;
} HeapItem;

static void heap_push(HeapItem* h, int64_t* n, double key, int32_t v) {
    int64_t i = (*n)++;
    while (i > 0) {
        int64_t p = (i - 1) >> 1;
        if (h[p].key <= key) break;
        h[i] = h[p];
        i = p;
    }
    h[i].key = key;
    h[i].v = v;
}
static HeapItem heap_pop(HeapItem* h, int64_t* n) {
    HeapItem top = h[0], last = h[--(*n)];
    int64_t i = 0;
    for (;;) {
        int64_t c = 2 * i + 1;
        if (c >= *n) break;
        if (c + 1 < *n && h[c + 1].key < h[c].key) ++c;
        if (last.key <= h[c].key) break;
        h[i] = h[c];
        i = c;
    }
    h[i] = last;
    return top;
}

int oracle_dijkstra(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n, int32_t source,
                    double* dist) {
    int64_t nnz = indptr[n], hn = 0;
    HeapItem* heap = (HeapItem*)malloc(sizeof(HeapItem) * (size_t)(nnz + 2));
    if (!heap) return -1;
    for (int64_t i = 0; i < n; ++i) dist[i] = INFINITY;
    dist[source] = 0.0;
    heap_push(heap, &hn, 0.0, source);
    while (hn > 0) {
        HeapItem it = heap_pop(heap, &hn);
        if (it.key > dist[it.v]) continue; /* stale entry (lazy deletion) */
        for (int64_t p = indptr[it.v]; p < indptr[it.v + 1]; ++p) {
            double cand = it.key + (double)data[p];
            int32_t v = indices[p];
            if (cand < dist[v]) {
                dist[v] = cand;
                heap_push(heap, &hn, cand, v);
            }
        }
    }
    free(heap);
    return 0;
}

/* ---- NumPy float32 pairwise summation (numpy/_core/src/umath/loops_utils.h.src, FLOAT_pairwise_sum) ---- */
float oracle_pairwise_sum_f32(const float* a, int64_t n) {
    if (n < 8) {
        float r = 0.f;
        for (int64_t i = 0; i < n; ++i) r += a[i];
        return r;
    }
    if (n <= 128) {
        volatile float r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int64_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] = r[j] + a[i + j];
        float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return oracle_pairwise_sum_f32(a, n2) + oracle_pairwise_sum_f32(a + n2, n - n2);
}

/* ---- one k-means++ draw (kmeans_optimized.py:47-61 + RandomState.choice): returns the index or -1 when
 * probs.sum()==0.  dmin: float32 distances to the nearest centre; u: the MT19937 double of the draw. ---- */
int64_t oracle_kpp_draw(const float* dmin, const uint8_t* is_center, int64_t n, double u, float* probs_sum_out) {
    float* p = (float*)malloc(sizeof(float) * (size_t)n);
    float maxf = 0.f;
    int any = 0;
    for (int64_t i = 0; i < n; ++i)
        if (isfinite(dmin[i])) {
            any = 1;
            if (dmin[i] > maxf) maxf = dmin[i];
        }
    for (int64_t i = 0; i < n; ++i) {
        float s = any ? (isfinite(dmin[i]) ? dmin[i] : maxf * 2.0f) : 1.0f;
        p[i] = is_center[i] ? 0.f : s * s;
    }
    float tot = oracle_pairwise_sum_f32(p, n);
    if (probs_sum_out) *probs_sum_out = tot;
    int64_t found = -1;
    if (tot > 0.f) {
        double last = 0.0;
        for (int64_t i = 0; i < n; ++i) last += (double)(p[i] / tot); /* sequential fp64 cumsum */
        double run = 0.0;
        found = n;
        for (int64_t i = 0; i < n; ++i) {
            run += (double)(p[i] / tot);
            if (run / last > u) { /* searchsorted(cdf/cdf[-1], u, side='right') */
                found = i;
                break;
            }
        }
    }
    free(p);
    return found;
}
