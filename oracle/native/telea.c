/* ORACLE — test infrastructure, never shipped, never on the product path.
 *
 * CPU restatement of the Telea fast-marching inpainting that the reference reaches through the
 * third-party call cv2.inpaint(..., cv2.INPAINT_TELEA) (Code/shape_ftp.py:652-666 on float32 images,
 * radius 3; Code/temperature_sensor.py:546-577 on uint8 images, radius 7 / 5).  OpenCV is not vendored
 * under /root/reference (README.md:63 is an unpinned `pip install opencv-python`; 4.13.0 in the build
 * container), so this file restates the published algorithm (A. Telea, "An image inpainting technique
 * based on the fast marching method", JGT 2004) with the processing order of OpenCV's implementation:
 * a one-pixel padded work frame, narrow band = 3x3-cross dilation of the hole, a stable (FIFO among
 * equal keys) priority list, the outward distance field over a (2r+1)-square ring first, then the inward
 * march that fills pixels.  Pinned bit-exactly against cv2.inpaint in tests/test_oracle_native.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KNOWN 0
#define BAND 1
#define INSIDE 2
#define CHANGE 3

/* ---- stable priority queue: binary heap ordered by (T, insertion sequence) ---- */
typedef struct { float T; unsigned seq; int i, j; } Elem;
typedef struct { Elem* a; int n, cap; unsigned seq; } PQ;

static int pq_less(const Elem* x, const Elem* y) { return x->T < y->T || (x->T == y->T && x->seq < y->seq); }
static void pq_init(PQ* q, int cap) { q->a = (Elem*)malloc(sizeof(Elem) * (size_t)(cap + 1)); q->n = 0; q->cap = cap; q->seq = 0; }
static void pq_push(PQ* q, int i, int j, float T) {
  if (q->n >= q->cap) { q->cap = q->cap * 2 + 16; q->a = (Elem*)realloc(q->a, sizeof(Elem) * (size_t)(q->cap + 1)); }
  Elem e = {T, q->seq++, i, j};
  int k = q->n++;
  while (k > 0) {
    int p = (k - 1) >> 1;
    if (!pq_less(&e, &q->a[p])) break;
    q->a[k] = q->a[p];
    k = p;
  }
  q->a[k] = e;
}
static int pq_pop(PQ* q, int* i, int* j) {
  if (q->n == 0) return 0;
  *i = q->a[0].i; *j = q->a[0].j;
  Elem e = q->a[--q->n];
  int k = 0;
  for (;;) {
    int c = 2 * k + 1;
    if (c >= q->n) break;
    if (c + 1 < q->n && pq_less(&q->a[c + 1], &q->a[c])) c++;
    if (!pq_less(&q->a[c], &e)) break;
    q->a[k] = q->a[c];
    k = c;
  }
  if (q->n > 0) q->a[k] = e;
  return 1;
}

static float fm_solve(int i1, int j1, int i2, int j2, const uint8_t* f, const float* t, int W) {
  double sol, a11 = t[(size_t)i1 * W + j1], a22 = t[(size_t)i2 * W + j2], m12 = a11 < a22 ? a11 : a22;
  if (f[(size_t)i1 * W + j1] != INSIDE) {
    if (f[(size_t)i2 * W + j2] != INSIDE) {
      if (fabs(a11 - a22) >= 1.0) sol = 1 + m12;
      else sol = (a11 + a22 + sqrt((double)(2 - (a11 - a22) * (a11 - a22)))) * 0.5;
    } else sol = 1 + a11;
  } else if (f[(size_t)i2 * W + j2] != INSIDE) sol = 1 + a22;
  else sol = 1 + m12;
  return (float)sol;
}
static float min4(float a, float b, float c, float d) {
  a = a < b ? a : b; c = c < d ? c : d;
  return a < c ? a : c;
}
static float fm_dist(int i, int j, const uint8_t* f, const float* t, int W) {
  return min4(fm_solve(i - 1, j, i, j - 1, f, t, W), fm_solve(i + 1, j, i, j - 1, f, t, W),
              fm_solve(i - 1, j, i, j + 1, f, t, W), fm_solve(i + 1, j, i, j + 1, f, t, W));
}

/* img: h x w (float32 when is_u8 == 0, uint8 otherwise); mask: h x w non-zero = hole; out: same type as img.
 * If t_out != NULL it receives the (h+2) x (w+2) distance field (debugging). */
static void telea(const void* img_, const uint8_t* mask_in, int h, int w, int range, void* out_, int is_u8, float* t_out) {
  const int H = h + 2, W = w + 2;
  const size_t n = (size_t)H * W;
  uint8_t* f = (uint8_t*)calloc(n, 1);
  uint8_t* mask = (uint8_t*)calloc(n, 1);
  uint8_t* band = (uint8_t*)calloc(n, 1);
  uint8_t* ring = (uint8_t*)calloc(n, 1);
  float* t = (float*)malloc(sizeof(float) * n);
  float* out = (float*)malloc(sizeof(float) * (size_t)h * w);     /* working image in float (u8 values are exact) */
  for (size_t k = 0; k < (size_t)h * w; k++) out[k] = is_u8 ? (float)((const uint8_t*)img_)[k] : ((const float*)img_)[k];
  for (size_t k = 0; k < n; k++) t[k] = 1.0e6f;
  int nmask = 0;
  for (int i = 0; i < h; i++)
    for (int j = 0; j < w; j++)
      if (mask_in[(size_t)i * w + j]) { mask[(size_t)(i + 1) * W + j + 1] = 1; nmask++; }
  /* band = dilate(mask, 3x3 cross) - mask, border row/col cleared */
  int nband = 0;
  for (int i = 1; i < H - 1; i++)
    for (int j = 1; j < W - 1; j++) {
      size_t p = (size_t)i * W + j;
      if (!mask[p] && (mask[p - 1] || mask[p + 1] || mask[p - W] || mask[p + W])) { band[p] = 1; nband++; }
    }
  PQ heap, outq;
  pq_init(&heap, nband + nmask + 16);
  for (size_t p = 0; p < n; p++) {
    if (band[p]) { pq_push(&heap, (int)(p / W), (int)(p % W), 0.f); f[p] = BAND; t[p] = 0.f; }
    else if (mask[p]) f[p] = INSIDE;
  }
  /* outward distances over the ring = dilate(mask, (2r+1) square) - mask - band */
  uint8_t* fo = ring;
  int nring = 0;
  for (int i = 0; i < H; i++)
    for (int j = 0; j < W; j++) {
      size_t p = (size_t)i * W + j;
      if (mask[p] || band[p]) continue;
      int hit = 0;
      for (int k = i - range; k <= i + range && !hit; k++) {
        if (k < 0 || k >= H) continue;
        for (int l = j - range; l <= j + range; l++)
          if (l >= 0 && l < W && mask[(size_t)k * W + l]) { hit = 1; break; }
      }
      if (hit) { fo[p] = INSIDE; nring++; }
    }
  pq_init(&outq, nband + nring + 16);
  for (size_t p = 0; p < n; p++)
    if (band[p]) pq_push(&outq, (int)(p / W), (int)(p % W), 0.f);
  {
    int ii, jj;
    while (pq_pop(&outq, &ii, &jj)) {
      fo[(size_t)ii * W + jj] = CHANGE;
      for (int q = 0; q < 4; q++) {
        int i = ii + (q == 0 ? -1 : q == 2 ? 1 : 0), j = jj + (q == 1 ? -1 : q == 3 ? 1 : 0);
        if (i <= 0 || j <= 0 || i > H - 1 || j > W - 1) continue;     /* rows/cols of the padded frame */
        if (i >= H - 1 || j >= W - 1) continue;
        if (fo[(size_t)i * W + j] == INSIDE) {
          float d = fm_dist(i, j, fo, t, W);
          t[(size_t)i * W + j] = d;
          fo[(size_t)i * W + j] = BAND;
          pq_push(&outq, i, j, d);
        }
      }
    }
    for (size_t p = 0; p < n; p++)
      if (fo[p] == CHANGE) { fo[p] = KNOWN; t[p] = -t[p]; }
  }
  /* inward march */
  {
    int ii, jj;
    while (pq_pop(&heap, &ii, &jj)) {
      f[(size_t)ii * W + jj] = KNOWN;
      for (int q = 0; q < 4; q++) {
        int i = ii + (q == 0 ? -1 : q == 2 ? 1 : 0), j = jj + (q == 1 ? -1 : q == 3 ? 1 : 0);
        if (i <= 0 || j <= 0 || i > H - 1 || j > W - 1) continue;
        if (i >= H - 1 || j >= W - 1) continue;
        const size_t p = (size_t)i * W + j;
        if (f[p] != INSIDE) continue;
        float dist = fm_dist(i, j, f, t, W);
        t[p] = dist;
        float gTx, gTy;
        if (f[p + 1] != INSIDE) {
          if (f[p - 1] != INSIDE) gTx = (t[p + 1] - t[p - 1]) * 0.5f; else gTx = t[p + 1] - t[p];
        } else {
          if (f[p - 1] != INSIDE) gTx = t[p] - t[p - 1]; else gTx = 0;
        }
        if (f[p + W] != INSIDE) {
          if (f[p - W] != INSIDE) gTy = (t[p + W] - t[p - W]) * 0.5f; else gTy = t[p + W] - t[p];
        } else {
          if (f[p - W] != INSIDE) gTy = t[p] - t[p - W]; else gTy = 0;
        }
        float Ia = 0, Jx = 0, Jy = 0, s = 1.0e-20f;
        for (int k = i - range; k <= i + range; k++) {
          int km = k - 1 + (k == 1), kp = k - 1 - (k == H - 2);
          for (int l = j - range; l <= j + range; l++) {
            int lm = l - 1 + (l == 1), lp = l - 1 - (l == W - 2);
            if (k > 0 && l > 0 && k < H - 1 && l < W - 1) {
              const size_t pk = (size_t)k * W + l;
              if (f[pk] != INSIDE && ((l - j) * (l - j) + (k - i) * (k - i) <= range * range)) {
                float ry = (float)(i - k), rx = (float)(j - l);
                float len2 = rx * rx + ry * ry;
                float dst = (float)(1. / (len2 * sqrtf(len2)));
                float lev = (float)(1. / (1 + fabs(t[pk] - t[p])));
                float dir = rx * gTx + ry * gTy;
                if (fabs(dir) <= 0.01) dir = 0.000001f;
                float wgt = (float)fabs(dst * lev * dir);
                float gIx, gIy;
#define O(a, b) out[(size_t)(a) * w + (b)]
                if (f[pk + 1] != INSIDE) {
                  if (f[pk - 1] != INSIDE) gIx = (O(km, lp + 1) - O(km, lm - 1)) * 2.0f; else gIx = O(km, lp + 1) - O(km, lm);
                } else {
                  if (f[pk - 1] != INSIDE) gIx = O(km, lp) - O(km, lm - 1); else gIx = 0;
                }
                if (f[pk + W] != INSIDE) {
                  if (f[pk - W] != INSIDE) gIy = (O(kp + 1, lm) - O(km - 1, lm)) * 2.0f; else gIy = O(kp + 1, lm) - O(km, lm);
                } else {
                  if (f[pk - W] != INSIDE) gIy = O(kp, lm) - O(km - 1, lm); else gIy = 0;
                }
                Ia += wgt * O(k - 1, l - 1);
                Jx -= wgt * (gIx * rx);
                Jy -= wgt * (gIy * ry);
                s += wgt;
              }
            }
          }
        }
        float sat;
        if (is_u8) {
          sat = (float)((Ia / s + (Jx + Jy) / (sqrtf(Jx * Jx + Jy * Jy) + 1.0e-20f) + 0.5f));
          int v = (int)lrintf(sat);                       /* cv::saturate_cast<uchar>(float) rounds to nearest even */
          v = v < 0 ? 0 : v > 255 ? 255 : v;
          O(i - 1, j - 1) = (float)v;
        } else {
          sat = (float)((Ia / s + (Jx + Jy) / (sqrtf(Jx * Jx + Jy * Jy) + 1.0e-20f)));
          O(i - 1, j - 1) = sat;
        }
#undef O
        f[p] = BAND;
        pq_push(&heap, i, j, dist);
      }
    }
  }
  if (is_u8) for (size_t k = 0; k < (size_t)h * w; k++) ((uint8_t*)out_)[k] = (uint8_t)out[k];
  else memcpy(out_, out, sizeof(float) * (size_t)h * w);
  if (t_out) memcpy(t_out, t, sizeof(float) * n);
  free(heap.a); free(outq.a); free(f); free(mask); free(band); free(ring); free(t); free(out);
}

void vo_inpaint_telea_f32(const float* img, const uint8_t* mask, int h, int w, int range, float* out, float* t_out) {
  telea(img, mask, h, w, range, out, 0, t_out);
}
void vo_inpaint_telea_u8(const uint8_t* img, const uint8_t* mask, int h, int w, int range, uint8_t* out, float* t_out) {
  telea(img, mask, h, w, range, out, 1, t_out);
}
