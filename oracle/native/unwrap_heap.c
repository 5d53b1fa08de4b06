/* ORACLE (test infrastructure, never shipped): CPU restatement of the reference's
 * quality-guided flood unwrap, shape_ftp.py:1043-1080 (unwrap_quality_guided).
 *
 * The reference keeps a Python heapq of tuples (-q, y, x, py, px); tuples compare
 * lexicographically, so the pop order is a total order and any correct priority
 * queue reproduces it.  Arithmetic follows NumPy >= 2 scalar semantics:
 *   dw = angle(exp(1j * (w[y,x] - w[py,px])))  evaluated in complex64/float32,
 *   u[y,x] = u[py,px] + dw                      in float32.
 */
#include <stdlib.h>
#include <stdint.h>
#include <math.h>
#include <complex.h>

typedef struct { double negq; int32_t y, x, py, px; } item_t;

static inline int item_less(const item_t* a, const item_t* b) {
    if (a->negq != b->negq) return a->negq < b->negq;
    if (a->y != b->y) return a->y < b->y;
    if (a->x != b->x) return a->x < b->x;
    if (a->py != b->py) return a->py < b->py;
    return a->px < b->px;
}

typedef struct { item_t* v; size_t n, cap; } heap_t;

static void heap_push(heap_t* h, item_t it) {
    if (h->n == h->cap) { h->cap = h->cap ? h->cap * 2 : (1u << 16); h->v = (item_t*)realloc(h->v, h->cap * sizeof(item_t)); }
    size_t i = h->n++;
    while (i > 0) {
        size_t p = (i - 1) >> 1;
        if (!item_less(&it, &h->v[p])) break;
        h->v[i] = h->v[p]; i = p;
    }
    h->v[i] = it;
}

static item_t heap_pop(heap_t* h) {
    item_t top = h->v[0];
    item_t last = h->v[--h->n];
    size_t i = 0, n = h->n;
    for (;;) {
        size_t c = 2 * i + 1;
        if (c >= n) break;
        if (c + 1 < n && item_less(&h->v[c + 1], &h->v[c])) c++;
        if (!item_less(&h->v[c], &last)) break;
        h->v[i] = h->v[c]; i = c;
    }
    if (n) h->v[i] = last;
    return top;
}

/* wrapped, quality: float32 [h*w]; mask: uint8 [h*w]; out: float32 [h*w] (NaN where unreached).
 * Returns number of pixels unwrapped. */
long vo_unwrap_quality_guided(const float* wrapped, const uint8_t* mask, const float* quality,
                              int h, int w, float* out, const float* dw8)
{
    const long n = (long)h * w;
    for (long i = 0; i < n; i++) out[i] = NAN;
    long seed = -1; float best = -INFINITY;
    int any = 0;
    for (long i = 0; i < n; i++) if (mask[i]) { any = 1; break; }
    if (!any) return 0;
    /* np.argmax over q with q[~m] = -inf: first maximum in raster order (NaN would win; not expected) */
    for (long i = 0; i < n; i++) {
        float q = mask[i] ? quality[i] : -INFINITY;
        if (seed < 0 || q > best) { best = q; seed = i; }
    }
    static const int dy8[8] = {-1, 1, 0, 0, -1, -1, 1, 1};
    static const int dx8[8] = { 0, 0,-1, 1, -1,  1,-1, 1};
    heap_t hp = {0, 0, 0};
    out[seed] = wrapped[seed];
    long done = 1;
    int sy = (int)(seed / w), sx = (int)(seed % w);
    #define PUSH_NB(PY, PX) do { \
        for (int k = 0; k < 8; k++) { int ny = (PY) + dy8[k], nx = (PX) + dx8[k]; \
            if (ny >= 0 && ny < h && nx >= 0 && nx < w) { long j = (long)ny * w + nx; \
                if (mask[j] && !isfinite(out[j])) { item_t nb_ = { -(double)quality[j], ny, nx, (PY), (PX) }; heap_push(&hp, nb_); } } } } while (0)
    PUSH_NB(sy, sx);
    while (hp.n) {
        item_t it = heap_pop(&hp);
        long j = (long)it.y * w + it.x, p = (long)it.py * w + it.px;
        if (!mask[j]) continue;
        if (isfinite(out[j])) continue;
        if (!isfinite(out[p])) continue;
        float dw;
        if (dw8) {
            /* dw8[k][j]: wrapped step from parent direction k into pixel j, precomputed by NumPy
             * (bit-identical to the reference's scalar np.angle(np.exp(1j*d))). */
            int k = 0;
            for (; k < 8; k++) if (it.y - dy8[k] == it.py && it.x - dx8[k] == it.px) break;
            dw = dw8[(long)k * n + j];
        } else {
            float d = wrapped[j] - wrapped[p];
            float complex e = cexpf(d * I);
            dw = atan2f(cimagf(e), crealf(e));
        }
        out[j] = out[p] + dw;
        done++;
        PUSH_NB(it.y, it.x);
    }
    free(hp.v);
    return done;
}
