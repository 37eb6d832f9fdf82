/*
 * rebvio_oracle.c — plain-C CPU restatement of the REBVIO hot path. TEST INFRASTRUCTURE (see rebvio_oracle.h).
 *
 * Every function cites the reference lines it restates (paths relative to the reference root). The code is
 * written for bit-equality with the reference built with `-O2 -ffp-contract=off`: float32 arithmetic in the
 * reference's evaluation order, `double` exactly where a double literal promotes the reference expression.
 * Parity is pinned by tests/test_oracle_vs_ref.py against oracle/_ref/librebvio_ref.so (the unmodified
 * reference sources) — see the header for the one tolerance-level exception (orc_solve6).
 */
#define _GNU_SOURCE
#include "rebvio_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define RHO_MAX 20.0f  /* types/keyline.hpp:19 */
#define RHO_MIN 1e-3f  /* types/keyline.hpp:20 */
#define RHO_INIT 1.0f  /* types/keyline.hpp:21 */
#define FIT 2          /* EdgeDetectorConfig::plane_fit_size, edge_detector.hpp:22 */
#define TUNE_BINS 100  /* EdgeDetectorConfig::num_bins, edge_detector.hpp:29 */

struct orc_map {
  rbv_keyline* kl;
  int n, cap;
  int* mask; /* pixel -> keyline id, -1 = none (unordered_map keylines_mask_, edge_map.hpp:130) */
  int npix;
  float threshold; /* -1 until tuned (edge_map.cpp:17) */
  unsigned matches;
};

typedef struct {
  int id, distance;
} field_cell; /* DistanceFieldElement, core.hpp:15-18 */

struct orc_ctx {
  rbv_camera cam;
  rbv_detector_cfg det;
  rbv_map_cfg mapc;
  rbv_core_cfg core;
  int widths[2][3];
  float sigma_true[2];
  float* div[2][3];
  float pinv[3][25];
  float *s0, *s1, *dog, *mag, *tmpa, *tmpb;
  int* det_mask; /* EdgeDetector::keylines_mask_ (edge_detector.hpp:86) */
  field_cell* field;
  orc_map* field_map; /* DistanceField::map_ (core.hpp:78) */
};

/* ------------------------------------------------------------------------------------------------------ */
/* small fixed-size algebra (TooN semantics)                                                               */

/* determinant by Gaussian elimination with partial pivoting: TooN determinant.h:91-146 (3x3 -> :186) */
static float det3(const float in[9]) {
  float A[3][3];
  for (int i = 0; i < 9; ++i) A[i / 3][i % 3] = in[i];
  float det = 1;
  for (int i = 0; i < 3; ++i) {
    int arg = i;
    float best = fabsf(A[i][i]);
    for (int k = i + 1; k < 3; ++k) {
      float v = fabsf(A[k][i]);
      if (v > best) { best = v; arg = k; }
    }
    float pivot = A[arg][i];
    if (arg != i) {
      det *= -1;
      for (int j = i; j < 3; ++j) { float t = A[i][j]; A[i][j] = A[arg][j]; A[arg][j] = t; }
    }
    det *= A[i][i];
    if (det == 0) return 0;
    for (int u = i + 1; u < 3; ++u) {
      float factor = A[u][i] / pivot;
      for (int j = i + 1; j < 3; ++j) A[u][j] = A[u][j] - factor * A[i][j];
    }
  }
  return det;
}

/* cofactor inverse divided by the determinant: types/definitions.hpp:40-53 */
static void inv3(const float m[9], float out[9]) {
  float co[9];
  co[0] = m[4] * m[8] - m[5] * m[7];
  co[1] = m[2] * m[7] - m[1] * m[8];
  co[2] = m[1] * m[5] - m[2] * m[4];
  co[3] = m[5] * m[6] - m[3] * m[8];
  co[4] = m[0] * m[8] - m[2] * m[6];
  co[5] = m[2] * m[3] - m[0] * m[5];
  co[6] = m[3] * m[7] - m[4] * m[6];
  co[7] = m[1] * m[6] - m[0] * m[7];
  co[8] = m[0] * m[4] - m[1] * m[3];
  float d = det3(m);
  for (int i = 0; i < 9; ++i) out[i] = co[i] / d;
}

/* ------------------------------------------------------------------------------------------------------ */
/* setup                                                                                                   */

/* FastGaussian::FastGaussian, scale_space.cpp:14-35 (Kovesi box widths) */
static void box_widths(float sigma, int n, int* widths, float* sigma_true) {
  float w_ideal = (float)sqrt(12.0 * sigma * sigma / (float)(n + 1));
  int wl = (int)w_ideal;
  if ((wl / 2) * 2 == wl) --wl;
  /* :25 — int terms, float sigma promoted by the int literals -> float arithmetic, std::round(float) */
  int m = (int)roundf(((float)(3 * n + 4 * n * wl + n * wl * wl) - 12 * sigma * sigma) / (float)(4 + 4 * wl));
  for (int i = 0; i < n; ++i) widths[i] = (i < m) ? wl : wl + 2;
  *sigma_true = (float)sqrt((m * wl * wl + (n - m) * (wl + 2.0) * (wl + 2.0) - n) / 12.0);
}

/* FastGaussian::precomputeDivisors, scale_space.cpp:130-171: reciprocal of the clipped window area */
static void fill_divisors(int rows, int cols, int d, float* div) {
  int h = d / 2;
  for (int r = 0; r < rows; ++r) {
    int nr = (r < h + 1) ? (r + h + 1) : (r >= rows - h ? rows - r + h : d);
    for (int c = 0; c < cols; ++c) {
      int nc = (c < h + 1) ? (c + h + 1) : (c >= cols - h ? cols - c + h : d);
      float area = (float)nr * (float)nc; /* exact small integers in every branch of :133-165 */
      div[r * cols + c] = (float)(1.0 / area); /* :168 double division, stored as float */
    }
  }
}

orc_ctx* orc_create(const rbv_camera* cam, const rbv_detector_cfg* det, const rbv_map_cfg* map, const rbv_core_cfg* core) {
  orc_ctx* c = (orc_ctx*)calloc(1, sizeof(orc_ctx));
  c->cam = *cam; c->det = *det; c->mapc = *map; c->core = *core;
  size_t n = (size_t)cam->rows * cam->cols;
  /* ScaleSpace::ScaleSpace, scale_space.cpp:186: sigma0 = 3.56359, sigma1 = sigma0_true * 1.2599 */
  box_widths(3.56359f, 3, c->widths[0], &c->sigma_true[0]);
  box_widths((float)(c->sigma_true[0] * 1.2599), 3, c->widths[1], &c->sigma_true[1]);
  for (int s = 0; s < 2; ++s)
    for (int p = 0; p < 3; ++p) {
      c->div[s][p] = (float*)malloc(n * sizeof(float));
      fill_divisors(cam->rows, cam->cols, c->widths[s][p], c->div[s][p]);
    }
  /* plane-fit pseudo inverse, edge_detector.cpp:55-68: Pinv = invert(Phi^T Phi) * Phi^T, Phi rows (col,row,1) */
  float phi[25][3], ptp[9], ipt[9];
  for (int r = -FIT, k = 0; r <= FIT; ++r)
    for (int q = -FIT; q <= FIT; ++q, ++k) { phi[k][0] = (float)q; phi[k][1] = (float)r; phi[k][2] = 1.0f; }
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) {
      float acc = 0;
      for (int k = 0; k < 25; ++k) acc += phi[k][a] * phi[k][b];
      ptp[a * 3 + b] = acc;
    }
  inv3(ptp, ipt);
  for (int a = 0; a < 3; ++a)
    for (int k = 0; k < 25; ++k) {
      float acc = 0;
      for (int b = 0; b < 3; ++b) acc += ipt[a * 3 + b] * phi[k][b];
      c->pinv[a][k] = acc;
    }
  c->s0 = (float*)calloc(n, sizeof(float)); c->s1 = (float*)calloc(n, sizeof(float));
  c->dog = (float*)calloc(n, sizeof(float)); c->mag = (float*)calloc(n, sizeof(float)); /* zero init :188-189 */
  c->tmpa = (float*)malloc(n * sizeof(float)); c->tmpb = (float*)malloc(n * sizeof(float));
  c->det_mask = (int*)malloc(n * sizeof(int));
  for (size_t i = 0; i < n; ++i) c->det_mask[i] = -1; /* edge_detector.cpp:25 */
  c->field = (field_cell*)malloc(n * sizeof(field_cell));
  for (size_t i = 0; i < n; ++i) { c->field[i].id = -1; c->field[i].distance = 2147483647; }
  return c;
}

void orc_destroy(orc_ctx* c) {
  if (!c) return;
  for (int s = 0; s < 2; ++s) for (int p = 0; p < 3; ++p) free(c->div[s][p]);
  free(c->s0); free(c->s1); free(c->dog); free(c->mag); free(c->tmpa); free(c->tmpb); free(c->det_mask); free(c->field);
  free(c);
}

void orc_scale_setup(orc_ctx* c, int* widths6, float* sigma_true2) {
  for (int s = 0; s < 2; ++s) { for (int p = 0; p < 3; ++p) widths6[s * 3 + p] = c->widths[s][p]; sigma_true2[s] = c->sigma_true[s]; }
}
void orc_divisors(orc_ctx* c, int scale, int pass, float* out) {
  memcpy(out, c->div[scale][pass], sizeof(float) * (size_t)c->cam.rows * c->cam.cols);
}
void orc_pinv(orc_ctx* c, float* pinv75) { memcpy(pinv75, c->pinv, sizeof(c->pinv)); }

/* ------------------------------------------------------------------------------------------------------ */
/* scale space                                                                                             */

/* FastGaussian::createIntegralImage, scale_space.cpp:48-67: sequential row prefix, then column prefix */
static void integral_image(int rows, int cols, const float* in, float* ii) {
  for (int r = 0; r < rows; ++r) {
    const float* src = in + (size_t)r * cols;
    float* dst = ii + (size_t)r * cols;
    float run = src[0];
    dst[0] = run;
    for (int c = 1; c < cols; ++c) { run = run + src[c]; dst[c] = run; }
  }
  for (int r = 1; r < rows; ++r) {
    const float* up = ii + (size_t)(r - 1) * cols;
    float* dst = ii + (size_t)r * cols;
    for (int c = 0; c < cols; ++c) dst[c] = dst[c] + up[c];
  }
}

/* FastGaussian::average, scale_space.cpp:69-128. Nine regions collapsed into clipped taps:
 *   A = ii[min(r+h,R-1)][min(c+h,C-1)], B = ii[.][c-h-1], Cc = ii[r-h-1][.], D = ii[r-h-1][c-h-1]
 *   missing taps (r-h-1 < 0 or c-h-1 < 0) drop out; association ((A-B)-Cc)+D except in the bottom rows
 *   (r >= R-h) where the reference writes ((A-Cc)-B)+D (:119,:124); the interior multiplies by the constant
 *   a = float(1.0/(d*d)) (:72,:99), every border region by the divisor image. */
static void box_average(int rows, int cols, const float* ii, int d, const float* div, float* out) {
  int h = d / 2;
  float a = (float)(1.0 / (d * d));
  for (int r = 0; r < rows; ++r) {
    int rlo = r - h - 1, rhi = r + h;
    int bottom = rhi >= rows;
    if (bottom) rhi = rows - 1;
    const float* hi = ii + (size_t)rhi * cols;
    const float* lo = rlo >= 0 ? ii + (size_t)rlo * cols : 0;
    for (int c = 0; c < cols; ++c) {
      int clo = c - h - 1, chi = c + h;
      int right = chi >= cols;
      if (right) chi = cols - 1;
      float v;
      if (!lo && clo < 0) v = hi[chi];
      else if (!lo) v = hi[chi] - hi[clo];
      else if (clo < 0) v = hi[chi] - lo[chi];
      else if (bottom) v = ((hi[chi] - lo[chi]) - hi[clo]) + lo[clo];
      else v = ((hi[chi] - hi[clo]) - lo[chi]) + lo[clo];
      int interior = lo && !bottom && clo >= 0 && !right;
      out[(size_t)r * cols + c] = v * (interior ? a : div[(size_t)r * cols + c]);
    }
  }
}

void orc_integral(orc_ctx* c, const float* in, float* out) { integral_image(c->cam.rows, c->cam.cols, in, out); }

void orc_box_pass(orc_ctx* c, int scale, int pass, const float* in, float* out) {
  integral_image(c->cam.rows, c->cam.cols, in, c->tmpa);
  box_average(c->cam.rows, c->cam.cols, c->tmpa, c->widths[scale][pass], c->div[scale][pass], out);
}

/* FastGaussian::smooth, scale_space.cpp:173-182: three box passes */
static void smooth(orc_ctx* c, int scale, const float* img, float* out) {
  orc_box_pass(c, scale, 0, img, c->tmpb);
  orc_box_pass(c, scale, 1, c->tmpb, out);
  memcpy(c->tmpb, out, sizeof(float) * (size_t)c->cam.rows * c->cam.cols);
  orc_box_pass(c, scale, 2, c->tmpb, out);
}

/* ScaleSpace::build + calculateDoG + calculateGradientMagnitude, scale_space.cpp:203-233 */
void orc_scale_space(orc_ctx* c, const float* img, float* s0, float* s1, float* dog, float* mag) {
  int rows = c->cam.rows, cols = c->cam.cols;
  size_t n = (size_t)rows * cols;
  smooth(c, 0, img, c->s0);
  smooth(c, 1, img, c->s1);
  for (size_t i = 0; i < n; ++i) c->dog[i] = c->s1[i] - c->s0[i];
  for (int r = 1; r < rows - 1; ++r)
    for (int q = 1; q < cols - 1; ++q) {
      float dx = c->s0[(size_t)r * cols + q + 1] - c->s0[(size_t)r * cols + q - 1];
      float dy = c->s0[(size_t)(r + 1) * cols + q] - c->s0[(size_t)(r - 1) * cols + q];
      c->mag[(size_t)r * cols + q] = dx * dx + dy * dy; /* border stays 0 */
    }
  if (s0) memcpy(s0, c->s0, n * sizeof(float));
  if (s1) memcpy(s1, c->s1, n * sizeof(float));
  if (dog) memcpy(dog, c->dog, n * sizeof(float));
  if (mag) memcpy(mag, c->mag, n * sizeof(float));
}

/* ------------------------------------------------------------------------------------------------------ */
/* maps                                                                                                    */

static orc_map* map_alloc(int cap, int npix) {
  orc_map* m = (orc_map*)calloc(1, sizeof(orc_map));
  m->cap = cap; m->npix = npix;
  m->kl = (rbv_keyline*)malloc(sizeof(rbv_keyline) * (size_t)(cap > 0 ? cap : 1));
  m->mask = (int*)malloc(sizeof(int) * (size_t)npix);
  for (int i = 0; i < npix; ++i) m->mask[i] = -1;
  m->threshold = -1.0f; /* edge_map.cpp:17 */
  return m;
}
orc_map* orc_map_new(orc_ctx* c) { return map_alloc(c->det.keylines_max, c->cam.rows * c->cam.cols); }
void orc_map_free(orc_map* m) { if (m) { free(m->kl); free(m->mask); free(m); } }
int orc_map_size(const orc_map* m) { return m->n; }
float orc_map_threshold(const orc_map* m) { return m->threshold; }
void orc_map_set_threshold(orc_map* m, float t) { m->threshold = t; }
void orc_map_get(const orc_map* m, rbv_keyline* out) { memcpy(out, m->kl, sizeof(rbv_keyline) * (size_t)m->n); }
void orc_map_set(orc_map* m, const rbv_keyline* in, const uint32_t* pixel_idx, int n) {
  if (n > m->cap) { m->kl = (rbv_keyline*)realloc(m->kl, sizeof(rbv_keyline) * (size_t)n); m->cap = n; }
  memcpy(m->kl, in, sizeof(rbv_keyline) * (size_t)n);
  m->n = n;
  if (pixel_idx) {
    for (int i = 0; i < m->npix; ++i) m->mask[i] = -1;
    for (int i = 0; i < n; ++i) if (m->mask[pixel_idx[i]] < 0) m->mask[pixel_idx[i]] = i; /* emplace keeps the first */
  }
}
void orc_map_pixels(const orc_map* m, uint32_t* out_pixel) {
  for (int i = 0; i < m->npix; ++i) if (m->mask[i] >= 0) out_pixel[m->mask[i]] = (uint32_t)i;
}

/* KeyLine::KeyLine, types/keyline.hpp:42-59 */
static void keyline_init(rbv_keyline* k, float px, float py, float gx, float gy, float ix, float iy) {
  k->pos[0] = px; k->pos[1] = py;
  k->pos_img[0] = ix; k->pos_img[1] = iy;
  k->match_pos_img[0] = ix; k->match_pos_img[1] = iy;
  k->gradient[0] = gx; k->gradient[1] = gy;
  k->match_gradient[0] = 0; k->match_gradient[1] = 0;
  k->gradient_norm = sqrtf(gx * gx + gy * gy);
  k->match_gradient_norm = 0;
  k->rho = RHO_INIT; k->sigma_rho = RHO_MAX; /* rho(1.0), sigma_rho(20.0) */
  k->id = k->id_prev = k->id_next = k->match_id = k->match_id_forward = k->match_id_keyframe = -1;
  k->matches = 0;
}

/* ------------------------------------------------------------------------------------------------------ */
/* front end                                                                                               */

/* threshold servo, edge_detector.cpp:33-36 */
float orc_servo_threshold(const rbv_detector_cfg* cfg, float threshold, int last_count) {
  if (cfg->gain > 0) {
    threshold -= cfg->gain * (float)(cfg->keylines_ref - last_count);
    threshold = threshold > cfg->max_threshold ? cfg->max_threshold : (threshold < cfg->min_threshold ? cfg->min_threshold : threshold);
  }
  return threshold;
}

/* EdgeDetector::buildEdgeMap per-pixel loop, edge_detector.cpp:69-121 */
static void build_edge_map(orc_ctx* c, float threshold, orc_map* map) {
  int rows = c->cam.rows, cols = c->cam.cols;
  const int max_image_value = 765; /* :21 */
  float pn_threshold = (float)((2.0 * FIT + 1.0) * (2.0 * FIT + 1.0)) * c->det.pos_neg_threshold;                      /* :69 */
  float gt = (threshold * max_image_value * c->det.dog_threshold);
  float gradient_threshold_squared = gt * gt;                                                                           /* :70 */
  float mag_threshold = (threshold * max_image_value) * (threshold * max_image_value);                                  /* :71 */
  map->n = 0;
  for (int i = 0; i < map->npix; ++i) map->mask[i] = -1;
  int count = 0;
  for (int r = FIT; r < rows - FIT; ++r) {
    for (int q = FIT; q < cols - FIT; ++q) {
      int idx = q + r * cols;
      c->det_mask[idx] = -1;
      if (c->mag[idx] < mag_threshold) continue;                                                                        /* :83 */
      int pn = 0;
      float y[25];
      for (int dr = -FIT, k = 0; dr <= FIT; ++dr)
        for (int dc = -FIT; dc <= FIT; ++dc, ++k) {
          float v = c->dog[(r + dr) * cols + q + dc];
          y[k] = v;
          pn = (v > 0.0) ? pn + 1 : pn - 1;                                                                             /* :92 */
        }
      if (fabs((double)pn) > pn_threshold) continue;                                                                    /* :96 */
      float th[3];
      for (int a = 0; a < 3; ++a) { /* :98, TooN operators.hh:309-324 + :200-208: sequential dot from 0 */
        float acc = 0;
        for (int k = 0; k < 25; ++k) acc += c->pinv[a][k] * y[k];
        th[a] = acc;
      }
      float tmp = th[2] / (th[0] * th[0] + th[1] * th[1]);                                                              /* :99 */
      float xs = -th[0] * tmp, ys = -th[1] * tmp;
      if (fabs(xs) > 0.5 || fabs(ys) > 0.5) continue;                                                                   /* :103 */
      if (th[0] * th[0] + th[1] * th[1] < gradient_threshold_squared) continue;                                         /* :107 */
      float px = (float)q + xs, py = (float)r + ys;                                                                     /* :109 */
      keyline_init(&map->kl[count], px, py, th[0], th[1], px - c->cam.cx, py - c->cam.cy);                              /* :110 */
      c->det_mask[idx] = count;
      map->mask[idx] = count;
      if (++count >= c->det.keylines_max) {                                                                             /* :113-119 */
        for (++idx; idx < rows * cols; ++idx) c->det_mask[idx] = -1;
        map->n = count;
        return;
      }
    }
  }
  map->n = count;
}

/* EdgeDetector::joinEdges + nextKeylineIdx, edge_detector.cpp:125-165 */
static void join_edges(orc_ctx* c, orc_map* map) {
  int cols = c->cam.cols;
  for (int idx = 0; idx < map->n; ++idx) {
    rbv_keyline* k = &map->kl[idx];
    int x = (int)(k->pos[0] + 0.5), y = (int)(k->pos[1] + 0.5);
    float tx = -k->gradient[1], ty = k->gradient[0];
    /* neighbour probe order by tangent quadrant: horizontal, vertical, diagonal */
    int sx = (ty > 0.0) ? ((tx > 0.0) ? 1 : -1) : ((tx < 0.0) ? -1 : 1);
    int sy = (ty > 0.0) ? 1 : -1;
    int hit = c->det_mask[y * cols + x + sx];
    if (hit < 0) hit = c->det_mask[(y + sy) * cols + x];
    if (hit < 0) hit = c->det_mask[(y + sy) * cols + x + sx];
    if (hit < 0) continue;
    map->kl[hit].id_prev = idx;
    k->id_next = hit;
  }
}

/* EdgeDetector::tuneThreshold, edge_detector.cpp:167-186. The cumulative scan at :183 starts at bin 1 and
 * can only stop early when the bins after bin 0 already hold keylines_max keylines, which never happens
 * (bin 0 always holds the maximum), so it ends at i == num_bins; restated with that loop kept literal
 * (histogram[num_bins] — one past the end in the reference — is never needed to decide the exit). */
static void tune_threshold(orc_ctx* c, orc_map* map) {
  if (map->n < 1) return; /* reference: undefined behaviour on an empty map (:168) */
  float mx = map->kl[0].gradient_norm, mn = mx;
  for (int i = 1; i < map->n; ++i) {
    float g = map->kl[i].gradient_norm;
    if (mx < g) mx = g;
    if (mn > g) mn = g;
  }
  int hist[TUNE_BINS + 1] = {0};
  for (int i = 0; i < map->n; ++i) {
    float f = TUNE_BINS * (mx - map->kl[i].gradient_norm) / (mx - mn);
    int b = (f != f || f >= 2147483648.0f || f < -2147483648.0f) ? (-2147483647 - 1) : (int)f; /* cvttss2si */
    b = b > TUNE_BINS - 1 ? TUNE_BINS - 1 : b;
    b = b < 0 ? 0 : b;
    ++hist[b];
  }
  int i = 0;
  for (int a = 0; i < TUNE_BINS && a < c->det.keylines_max; i++, a += hist[i]) {}
  float thr = mx - (float)(i * (mx - mn)) / (float)TUNE_BINS;
  map->threshold = thr;
}

/* EdgeDetector::detect after the servo, edge_detector.cpp:37-39 */
int orc_detect(orc_ctx* c, const float* img, float threshold, orc_map* out) {
  orc_scale_space(c, img, 0, 0, 0, 0);
  out->threshold = -1.0f;
  build_edge_map(c, threshold, out);
  join_edges(c, out);
  tune_threshold(c, out);
  return out->n;
}

void orc_detector_mask(orc_ctx* c, int* out) { memcpy(out, c->det_mask, sizeof(int) * (size_t)c->cam.rows * c->cam.cols); }

/* ------------------------------------------------------------------------------------------------------ */
/* tracker                                                                                                 */

/* getIndex of core.hpp:66-71 / edge_map.hpp:119-124: std::round (half away from zero) then bounds test */
static int pixel_index(int rows, int cols, float row, float col) {
  int r = (int)roundf(row), q = (int)roundf(col);
  if (r >= rows || r < 0 || q >= cols || q < 0) return -1;
  return r * cols + q;
}

/* DistanceField::build, core.hpp:37-59 */
void orc_build_distance_field(orc_ctx* c, orc_map* m) {
  int rows = c->cam.rows, cols = c->cam.cols, n = rows * cols;
  c->field_map = m;
  for (int i = 0; i < n; ++i) c->field[i].id = -1; /* distance left stale, :41 */
  for (int idx = 0; idx < m->n; ++idx) {
    const rbv_keyline* k = &m->kl[idx];
    if (m->threshold > 0.0 && k->gradient_norm < m->threshold) continue;
    for (int r = (int)-c->core.search_range; r < c->core.search_range; ++r) {
      int fi = pixel_index(rows, cols, (k->gradient[1] / k->gradient_norm) * (float)r + k->pos[1],
                           (k->gradient[0] / k->gradient_norm) * (float)r + k->pos[0]);
      if (fi < 0) continue;
      field_cell* cell = &c->field[fi];
      int ar = r < 0 ? -r : r;
      if (cell->id >= 0 && cell->distance < ar) continue; /* ties go to the later (idx, r) */
      cell->distance = ar;
      cell->id = idx;
    }
  }
}

void orc_field(orc_ctx* c, int* id, int* distance) {
  int n = c->cam.rows * c->cam.cols;
  for (int i = 0; i < n; ++i) { id[i] = c->field[i].id; distance[i] = c->field[i].distance; }
}

/* EdgeMap::rotateKeylines, edge_map.cpp:58-71. makeVector(float,float,double) resolves to the
 * double overload (TooN make_vector.hh:26), so R*(..) is a float x double product accumulated in double
 * (operators.hh:200-208) and rounded to float on assignment to q. */
void orc_rotate(orc_ctx* c, orc_map* m, const float* R) {
  float fm = c->cam.fm;
  for (int i = 0; i < m->n; ++i) {
    rbv_keyline* k = &m->kl[i];
    double v[3] = {(double)(k->pos_img[0] / fm), (double)(k->pos_img[1] / fm), 1.0};
    float q[3];
    for (int r = 0; r < 3; ++r) {
      double acc = 0;
      for (int j = 0; j < 3; ++j) acc += R[r * 3 + j] * v[j];
      q[r] = (float)acc;
    }
    if (fabs(q[2]) > 0.0) {
      k->pos_img[0] = q[0] / q[2] * fm;
      k->pos_img[1] = q[1] / q[2] * fm;
      k->rho /= q[2];
      k->sigma_rho /= q[2];
    }
    double g[3] = {(double)k->gradient[0], (double)k->gradient[1], 0.0};
    for (int r = 0; r < 2; ++r) {
      double acc = 0;
      for (int j = 0; j < 3; ++j) acc += R[r * 3 + j] * g[j];
      q[r] = (float)acc;
    }
    k->gradient[0] = q[0];
    k->gradient[1] = q[1];
  }
}

static int double_to_int_x86(double d) { /* cvttsd2si: out-of-range / NaN -> INT_MIN */
  return (d != d || d >= 2147483648.0 || d <= -2147483649.0) ? (-2147483647 - 1) : (int)d;
}
static int float_to_int_x86(float f) { /* cvttss2si: out-of-range / NaN -> INT_MIN */
  return (f != f || f >= 2147483648.0f || f < -2147483648.0f) ? (-2147483647 - 1) : (int)f;
}

/* EdgeMap::estimateQuantile, edge_map.cpp:39-56 */
float orc_quantile(orc_map* m, float percentile, int num_bins) {
  int* hist = (int*)calloc((size_t)num_bins, sizeof(int));
  for (int i = 0; i < m->n; ++i) {
    int b = float_to_int_x86(num_bins * (m->kl[i].sigma_rho - RHO_MIN) / (RHO_MAX - RHO_MIN));
    b = b > num_bins - 1 ? num_bins - 1 : b;
    b = b < 0 ? 0 : b;
    ++hist[b];
  }
  float out = 1e3f;
  for (int i = 0, a = 0; i < num_bins; ++i) {
    if (a > percentile * m->n) { out = (float)i * (RHO_MAX - RHO_MIN) / (float)num_bins + RHO_MIN; break; }
    a += hist[i];
  }
  free(hist);
  return out;
}

/* Core::tryVel + calculatefJ + testfk, core.cpp:39-148.
 *
 * UNINITIALISED READ IN THE REFERENCE, MADE DETERMINISTIC: `fi` (core.cpp:122) is only assigned on the matched path
 * of calculatefJ (:67) but `_residuals[idx] = std::fabs(fi)` (:141) runs for every in-image keyline. In the compiled
 * reference the stack slot keeps the value of the most recent MATCHED keyline of the same call; before the first match
 * of a call it holds stack garbage (observed: the previous call's last |fi|, pointer halves, LM temporaries). The
 * restatement carries the last matched fi in index order and uses 0 before the first match of every call, which is
 * bit-identical to the reference run with a zero-filled stack (oracle/ref_harness.cpp: scrub_stack before each direct
 * Core::tryVel call) -- that is how tests/test_oracle_vs_ref.py pins it. */
float orc_try_vel(orc_ctx* c, orc_map* m, const float* vel, float sigma_rho_min, float* residuals, float* JtJ9, float* JtF3) {
  const float fm = c->cam.fm, cx = c->cam.cx, cy = c->cam.cy, range = c->core.search_range, rw = c->core.reweight_distance;
  const int rows = c->cam.rows, cols = c->cam.cols;
  const orc_map* tgt = c->field_map;
  float score = 0, a00 = 0, a11 = 0, a22 = 0, a01 = 0, a02 = 0, a12 = 0, b0 = 0, b1 = 0, b2 = 0;
  /* std::min(config_->min_match_threshold, frame_count_) with frame_count_ == 0 forever (core.cpp:22,91) */
  const unsigned min_matches = 0;
  float fi = 0; /* carried across iterations, see above */
  for (int idx = 0; idx < m->n; ++idx) {
    rbv_keyline* k = &m->kl[idx];
    k->match_id_forward = -1;
    if (m->threshold > 0.0 && k->gradient_norm < m->threshold) continue;
    if (k->sigma_rho > sigma_rho_min || k->matches < min_matches) continue;
    float w = 1.0;
    if (residuals[idx] > rw) w = rw / residuals[idx];
    float z_p = (float)(1.0 / k->rho + vel[2]);
    float f;
    if (z_p <= 0.0) {
      f = (float)((1.0 / k->sigma_rho) * range * w);
      score += f * f;
      continue;
    }
    float rho_p = (float)(1.0 / z_p);
    float px = rho_p * (vel[0] * fm - vel[2] * k->pos_img[0]) + k->pos_img[0];
    float py = rho_p * (vel[1] * fm - vel[2] * k->pos_img[1]) + k->pos_img[1];
    float pxc = px + cx, pyc = py + cy;
    int x = double_to_int_x86(pxc + 0.5), y = double_to_int_x86(pyc + 0.5);
    if (x < 1 || y < 1 || x >= cols - 1 || y >= rows - 1) {
      f = (float)((1.0 / k->sigma_rho) * range * w);
      score += f * f;
      continue;
    }
    float dfx = 0, dfy = 0;
    int tid = c->field[y * cols + x].id;
    int matched = 0;
    if (tid >= 0) {
      const rbv_keyline* t = &tgt->kl[tid];
      float n2 = k->gradient_norm * k->gradient_norm; /* testfk(_keyline1 = t, _keyline2 = k), :39-44 */
      float dp = t->gradient[0] * k->gradient[0] + t->gradient[1] * k->gradient[1];
      if (!(fabsf(dp - n2) > c->core.match_treshold * n2)) {
        matched = 1;
        float dx = pxc - t->pos[0], dy = pyc - t->pos[1];
        float gnx = t->gradient[0] / t->gradient_norm, gny = t->gradient[1] / t->gradient_norm;
        fi = dx * gnx + dy * gny;
        dfx = gnx / k->sigma_rho;
        dfy = gny / k->sigma_rho;
        k->match_id_forward = tid;
        f = fi / k->sigma_rho;
      }
    }
    if (!matched) f = range / k->sigma_rho;
    f *= w;
    score += f * f;
    float jx = rho_p * fm * dfx * w;
    float jy = rho_p * fm * dfy * w;
    float jz = -rho_p * (px * dfx + py * dfy) * w;
    a00 += jx * jx; a11 += jy * jy; a22 += jz * jz;
    a01 += jx * jy; a02 += jx * jz; a12 += jy * jz;
    b0 += jx * f; b1 += jy * f; b2 += jz * f;
    residuals[idx] = fabsf(fi);
  }
  JtJ9[0] = a00; JtJ9[4] = a11; JtJ9[8] = a22;
  JtJ9[1] = JtJ9[3] = a01; JtJ9[2] = JtJ9[6] = a02; JtJ9[5] = JtJ9[7] = a12;
  JtF3[0] = b0; JtF3[1] = b1; JtF3[2] = b2;
  return score;
}

/* Core::minimizeVel, core.cpp:150-189. trace (optional): per tryVel call t = 0..iterations, 16 floats:
 * vel[3], JtJ[9], JtF[3], score. */
float orc_minimize_vel(orc_ctx* c, orc_map* m, float* vel, float* Rvel9, float* trace, float* residuals_out) {
  float sigma_rho_min = orc_quantile(m, c->core.quantile_cutoff, c->core.quantile_num_bins);
  float* residuals = (float*)calloc((size_t)(m->n > 0 ? m->n : 1), sizeof(float));
  float JtJ[9], JtF[3], JtJn[9], JtFn[3], ApI[9], inv[9], h[3], vnew[3];
  float F = orc_try_vel(c, m, vel, sigma_rho_min, residuals, JtJ, JtF);
  if (trace) { memcpy(trace, vel, 12); memcpy(trace + 3, JtJ, 36); memcpy(trace + 12, JtF, 12); trace[15] = F; }
  float v = 2.0f, tau = 1e-3f;
  float mx = JtJ[0];
  for (int i = 1; i < 9; ++i) if (JtJ[i] > mx) mx = JtJ[i]; /* TooN::max_element, helpers.h:831 */
  float u = tau * mx;
  for (unsigned it = 0; it < c->core.iterations; ++it) {
    memcpy(ApI, JtJ, sizeof(ApI));
    ApI[0] = JtJ[0] + u; ApI[4] = JtJ[4] + u; ApI[8] = JtJ[8] + u;          /* JtJ + Identity*u, :167 */
    inv3(ApI, inv);
    float g[3] = {JtF[0] * -1, JtF[1] * -1, JtF[2] * -1};                   /* -JtF, operators.hh:187 */
    for (int r = 0; r < 3; ++r) {                                            /* h = inv * (-JtF), :168 */
      float acc = 0;
      for (int j = 0; j < 3; ++j) acc += inv[r * 3 + j] * g[j];
      h[r] = acc;
    }
    for (int i = 0; i < 3; ++i) vnew[i] = vel[i] + h[i];
    float Fn = orc_try_vel(c, m, vnew, sigma_rho_min, residuals, JtJn, JtFn);
    if (trace) { float* t = trace + (it + 1) * 16; memcpy(t, vnew, 12); memcpy(t + 3, JtJn, 36); memcpy(t + 12, JtFn, 12); t[15] = Fn; }
    double den = 0;                                                           /* 0.5*h * (u*h - JtF), :172 */
    for (int i = 0; i < 3; ++i) { float d = u * h[i] - JtF[i]; den += (0.5 * h[i]) * d; }
    float gain = (float)((F - Fn) / den);
    if (gain > 0.0) {
      F = Fn;
      memcpy(vel, vnew, 12); memcpy(JtJ, JtJn, 36); memcpy(JtF, JtFn, 12);
      double t = 2.0 * gain - 1.0;
      double s = 1.0 - (t * t * t);
      u = (float)(u * (0.33 > s ? 0.33 : s));                                /* std::max(0.33, ..), :178 */
      v = 2.0f;
    } else {
      u *= v;
      v = (float)(v * 2.0);
    }
  }
  inv3(JtJ, Rvel9);                                                          /* :185 */
  if (residuals_out) memcpy(residuals_out, residuals, sizeof(float) * (size_t)m->n);
  free(residuals);
  return F;
}

/* EdgeMap::forwardMatch, edge_map.cpp:73-99 */
int orc_forward_match(orc_map* old_map, orc_map* new_map) {
  unsigned n = 0;
  for (int idx = 0; idx < old_map->n; ++idx) {
    const rbv_keyline* k = &old_map->kl[idx];
    int f = k->match_id_forward;
    if (f < 0) continue;
    rbv_keyline* t = &new_map->kl[f];
    if (t->match_id >= 0 && t->rho > k->rho) continue;
    t->rho = k->rho;
    t->sigma_rho = k->sigma_rho;
    t->matches = k->matches + 1;
    t->match_id = idx;
    t->match_pos_img[0] = k->pos_img[0]; t->match_pos_img[1] = k->pos_img[1];
    t->match_gradient[0] = k->gradient[0]; t->match_gradient[1] = k->gradient[1];
    t->match_gradient_norm = k->gradient_norm;
    t->match_id_keyframe = k->match_id_keyframe;
    ++n;
  }
  new_map->matches = n;
  return (int)n;
}

/* symmetric 6x6 pseudo-inverse solve standing in for TooN::SVD<6,6,float>(JtJ).backsub(JtF) (core.cpp:247-248,
 * TooN SVD.h:122-163 -> LAPACK sgesvd_, un-vendored). Cyclic Jacobi eigen-decomposition in double; singular
 * values |lambda_i|; inverse zeroed when s_i * 1e9 <= s_max (SVD.h:37,266-274). Tolerance-level parity only. */
void orc_solve6(const float* A36, const float* b6, float* x6) {
  double A[6][6], V[6][6];
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { A[i][j] = 0.5 * ((double)A36[i * 6 + j] + (double)A36[j * 6 + i]); V[i][j] = i == j; }
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0;
    for (int p = 0; p < 6; ++p) for (int q = p + 1; q < 6; ++q) off += A[p][q] * A[p][q];
    if (off == 0) break;
    for (int p = 0; p < 6; ++p)
      for (int q = p + 1; q < 6; ++q) {
        if (A[p][q] == 0) continue;
        double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
        double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int k = 0; k < 6; ++k) { double akp = A[k][p], akq = A[k][q]; A[k][p] = cs * akp - sn * akq; A[k][q] = sn * akp + cs * akq; }
        for (int k = 0; k < 6; ++k) { double apk = A[p][k], aqk = A[q][k]; A[p][k] = cs * apk - sn * aqk; A[q][k] = sn * apk + cs * aqk; }
        for (int k = 0; k < 6; ++k) { double vkp = V[k][p], vkq = V[k][q]; V[k][p] = cs * vkp - sn * vkq; V[k][q] = sn * vkp + cs * vkq; }
      }
  }
  double smax = 0;
  for (int i = 0; i < 6; ++i) if (fabs(A[i][i]) > smax) smax = fabs(A[i][i]);
  double x[6] = {0};
  for (int i = 0; i < 6; ++i) {
    double s = fabs(A[i][i]);
    if (s * 1e9 <= smax) continue;
    double proj = 0;
    for (int k = 0; k < 6; ++k) proj += V[k][i] * (double)b6[k];
    proj /= A[i][i];
    for (int k = 0; k < 6; ++k) x[k] += V[k][i] * proj;
  }
  for (int k = 0; k < 6; ++k) x6[k] = (float)x[k];
}

/* Core::extRotVel, core.cpp:191-261 (on the distance-field map; `new_map` must be that map) */
int orc_ext_rot_vel(orc_ctx* c, orc_map* new_map, const float* vel, float* Wx36, float* X6, float* JtF6) {
  orc_map* m = c->field_map ? c->field_map : new_map;
  const float fm = c->cam.fm, pu = c->core.pixel_uncertainty, rw = c->core.reweight_distance;
  float JtJ[6][6] = {{0}}, JtF[6] = {0};
  for (int idx = 0; idx < m->n; ++idx) {
    const rbv_keyline* k = &m->kl[idx];
    if (k->match_id < 0) continue;
    float ux = k->gradient[0] / k->gradient_norm, uy = k->gradient[1] / k->gradient_norm;
    float rho_t = (float)(1.0 / (1.0 / k->rho + vel[2]));
    float mx = k->match_pos_img[0], my = k->match_pos_img[1];
    float qtx = mx + rho_t * (vel[0] * fm - vel[2] * mx);
    float qty = my + rho_t * (vel[1] * fm - vel[2] * my);
    float qx = k->pos_img[0], qy = k->pos_img[1];
    float phi[6];
    phi[0] = ux * rho_t * fm;
    phi[1] = uy * rho_t * fm;
    phi[2] = ux * (-rho_t * qx) + uy * (-rho_t * qy);
    phi[3] = -ux * qx * qy / fm - uy * (fm + qy * qy / fm);
    phi[4] = uy * qx * qy / fm + ux * (fm + qx * qx / fm);
    phi[5] = -ux * qy + uy * qx;
    float y = ux * (qx - qtx) + uy * (qy - qty);
    float dqvel = ux * (vel[0] * fm - vel[2] * mx) + uy * (vel[1] * fm - vel[2] * my);
    float sy = sqrtf(k->sigma_rho * k->sigma_rho * dqvel * dqvel + pu * pu);
    float w = 1.0f;
    if (fabsf(y) > rw) w = fabsf(y) / rw;
    float dv = sy * w;
    for (int a = 0; a < 6; ++a) phi[a] /= dv;
    y /= dv;
    /* JtJ = Phi^T Phi, JtF = Phi^T Y accumulated row by row == sequential dot over j (operators.hh:200-208) */
    for (int a = 0; a < 6; ++a) {
      for (int b = 0; b < 6; ++b) JtJ[a][b] += phi[a] * phi[b];
      JtF[a] += phi[a] * y;
    }
  }
  for (int a = 0; a < 6; ++a) for (int b = 0; b < 6; ++b) Wx36[a * 6 + b] = JtJ[a][b];
  if (JtF6) memcpy(JtF6, JtF, sizeof(JtF));
  orc_solve6(Wx36, JtF, X6);
  for (int i = 0; i < 6; ++i) if (X6[i] != X6[i]) return 0;
  return 1;
}

/* EdgeMap::searchMatch, edge_map.cpp:101-184: `self` is the map searched (old), `k` the querying keyline */
static int search_match(const orc_ctx* c, const orc_map* self, const rbv_keyline* k, const float* vel, const float* Rvel,
                        const float* Rback, float max_radius, unsigned long long* probes) {
  const float fm = c->cam.fm, cx = c->cam.cx, cy = c->cam.cy, pum = c->mapc.pixel_uncertainty_match;
  const float cang_min = (float)cos(c->mapc.match_threshold_angle * M_PI / 180.0);
  float in[3] = {k->pos_img[0], k->pos_img[1], fm}, pm[3];
  for (int r = 0; r < 3; ++r) { float acc = 0; for (int j = 0; j < 3; ++j) acc += Rback[r * 3 + j] * in[j]; pm[r] = acc; }
  float pmx = pm[0] * fm / pm[2], pmy = pm[1] * fm / pm[2];
  float k_rho = k->rho * fm / pm[2];
  float pi0x = pmx + cx, pi0y = pmy + cy;
  float tx = -(vel[0] * fm - vel[2] * pmx), ty = -(vel[1] * fm - vel[2] * pmy);
  float norm_t = sqrtf(tx * tx + ty * ty);
  float dr[3] = {fm, fm, -(pmx + pmy)}, row[3];
  for (int q = 0; q < 3; ++q) { float acc = 0; for (int j = 0; j < 3; ++j) acc += dr[j] * Rvel[j * 3 + q]; row[q] = acc; }
  float sigma2_t = 0;
  for (int j = 0; j < 3; ++j) sigma2_t += row[j] * dr[j];
  float dq_min, dq_max, dq_rho;
  int t_steps;
  if (norm_t > 1e-6) {
    tx /= norm_t; ty /= norm_t;
    dq_rho = norm_t * k_rho;
    float lo = norm_t * (k_rho - k->sigma_rho), hi = norm_t * (k_rho + k->sigma_rho);
    dq_min = (0.0f < lo ? lo : 0.0f) - pum;          /* std::max(0, lo) */
    dq_max = (hi < max_radius ? hi : max_radius) + pum; /* std::min(max_radius, hi) */
    if (dq_rho > dq_max) {
      dq_rho = (float)(0.5 * (dq_max + dq_min));
      t_steps = double_to_int_x86(dq_rho + 0.5);
    } else {
      float a = dq_max - dq_rho, b = dq_rho - dq_min;
      t_steps = float_to_int_x86(a < b ? b : a);
    }
  } else {
    tx = k->gradient[0]; ty = k->gradient[1];
    norm_t = k->gradient_norm;
    tx /= norm_t; ty /= norm_t;
    norm_t = 1.0f;
    dq_min = -max_radius - pum;
    dq_max = max_radius + pum;
    dq_rho = 0.0f;
    t_steps = float_to_int_x86(dq_max);
  }
  float tn = dq_rho, tp = dq_rho + 1;
  for (int ti = 0; ti < t_steps; ++ti, ++tp, --tn) {
    for (int side = 0; side < 2; ++side) {
      float t;
      if (side) { t = tp; if (t > dq_max) continue; }
      else { t = tn; if (t < dq_min) continue; }
      int idx = pixel_index(c->cam.rows, c->cam.cols, ty * t + pi0y, tx * t + pi0x);
      if (idx < 0) continue;
      if (probes) ++*probes;
      int id = self->mask[idx];
      if (id < 0) continue;
      const rbv_keyline* o = &self->kl[id];
      float cang = (o->gradient[0] * k->gradient[0] + o->gradient[1] * k->gradient[1]) / (o->gradient_norm * k->gradient_norm);
      if (cang < cang_min || fabs(o->gradient_norm / k->gradient_norm - 1.0) > c->mapc.match_threshold_norm) continue;
      float v_rho_dr = (pum * pum + o->sigma_rho * o->sigma_rho * norm_t * norm_t + sigma2_t * o->rho * o->rho);
      if ((t - norm_t * o->rho) * (t - norm_t * o->rho) > v_rho_dr) continue;
      return id;
    }
  }
  return -1;
}

/* EdgeMap::directedMatch, edge_map.cpp:186-218 (called on the new map, searching the old one) */
int orc_directed_match(orc_ctx* c, orc_map* new_map, orc_map* old_map, const float* vel_in, const float* Rvel_in,
                       const float* Rback, float max_radius, int* kf_matches, unsigned long long* probes) {
  float vel[3], tmp[9], Rvel[9];
  for (int r = 0; r < 3; ++r) { float acc = 0; for (int j = 0; j < 3; ++j) acc += Rback[r * 3 + j] * vel_in[j]; vel[r] = acc; }
  for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) { float acc = 0; for (int j = 0; j < 3; ++j) acc += Rback[r * 3 + j] * Rvel_in[j * 3 + q]; tmp[r * 3 + q] = acc; }
  for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) { float acc = 0; for (int j = 0; j < 3; ++j) acc += tmp[r * 3 + j] * Rback[q * 3 + j]; Rvel[r * 3 + q] = acc; }
  int n = 0, kf = 0;
  if (probes) *probes = 0;
  for (int idx = 0; idx < new_map->n; ++idx) {
    rbv_keyline* k = &new_map->kl[idx];
    int hit = search_match(c, old_map, k, vel, Rvel, Rback, max_radius, probes);
    if (hit < 0) continue;
    const rbv_keyline* o = &old_map->kl[hit];
    k->rho = o->rho;
    k->sigma_rho = o->sigma_rho;
    k->match_id = hit;
    k->matches = o->matches + 1;
    k->match_pos_img[0] = o->pos_img[0]; k->match_pos_img[1] = o->pos_img[1];
    k->match_gradient[0] = o->gradient[0]; k->match_gradient[1] = o->gradient[1];
    k->match_gradient_norm = o->gradient_norm;
    k->match_id_keyframe = o->match_id_keyframe;
    if (k->match_id_keyframe >= 0) ++kf;
    ++n;
  }
  new_map->matches = (unsigned)n;
  if (kf_matches) *kf_matches = kf;
  return n;
}

/* EdgeMap::regularize1Iter, edge_map.cpp:220-259 */
int orc_regularize(orc_ctx* c, orc_map* m) {
  const float thr = c->mapc.regularization_threshold;
  int n = m->n, count = 0;
  float* r = (float*)malloc(sizeof(float) * (size_t)(n + 1));
  float* s = (float*)malloc(sizeof(float) * (size_t)(n + 1));
  char* set = (char*)calloc((size_t)(n + 1), 1);
  for (int idx = 0; idx < n; ++idx) {
    const rbv_keyline* k = &m->kl[idx];
    if (k->id_next < 0 || k->id_prev < 0) continue;
    const rbv_keyline* kn = &m->kl[k->id_next];
    const rbv_keyline* kp = &m->kl[k->id_prev];
    if ((kn->rho - kp->rho) * (kn->rho - kp->rho) > (kn->sigma_rho * kn->sigma_rho + kp->sigma_rho * kp->sigma_rho)) continue;
    float alpha = (kn->gradient[0] * kp->gradient[0] + kn->gradient[1] * kp->gradient[1]) / (kn->gradient_norm * kp->gradient_norm);
    if (alpha < thr) continue;
    alpha = (float)((alpha - thr) / (1.0 - thr));
    alpha = (float)(alpha / (fabsf(kn->rho - kp->rho) / (kn->sigma_rho + kp->sigma_rho) + 1.0));
    float wr = (float)(1.0 / (k->sigma_rho * k->sigma_rho));
    float wrn = alpha / (kn->sigma_rho * kn->sigma_rho);
    float wrp = alpha / (kp->sigma_rho * kp->sigma_rho);
    r[idx] = (k->rho * wr + kn->rho * wrn + kp->rho * wrp) / (wr + wrn + wrp);
    s[idx] = (k->sigma_rho * wr + kn->sigma_rho * wrn + kp->sigma_rho * wrp) / (wr + wrn + wrp);
    set[idx] = 1;
    ++count;
  }
  for (int idx = 0; idx < n; ++idx) if (set[idx]) { m->kl[idx].rho = r[idx]; m->kl[idx].sigma_rho = s[idx]; }
  free(r); free(s); free(set);
  return count;
}

/* Core::updateInverseDepth + updateInverseDepthARLU, core.cpp:417-456 (on the distance-field map) */
void orc_update_inverse_depth(orc_ctx* c, const float* vel) {
  orc_map* m = c->field_map;
  if (!m) return;
  const float fm = c->cam.fm, pu = c->core.pixel_uncertainty, qa = c->core.reshape_q_abs;
  for (int idx = 0; idx < m->n; ++idx) {
    rbv_keyline* k = &m->kl[idx];
    if (k->match_id < 0) continue;
    float qx = k->pos_img[0], qy = k->pos_img[1], q0x = k->match_pos_img[0], q0y = k->match_pos_img[1];
    float v_rho = k->sigma_rho * k->sigma_rho;
    float ux = k->match_gradient[0] / k->match_gradient_norm, uy = k->match_gradient[1] / k->match_gradient_norm;
    float Y = ux * (qx - q0x) + uy * (qy - q0y);
    float H = ux * (vel[0] * fm - vel[2] * q0x) + uy * (vel[1] * fm - vel[2] * q0y);
    float rho_p = (float)(1.0 / (1.0 / k->rho + vel[2]));
    float F = (float)(1.0 / (1.0 + k->rho * vel[2]));
    F *= F;
    float p_p = F * v_rho * F + qa * qa;
    float e = Y - H * rho_p;
    float S = H * p_p * H + pu * pu;
    float K = (float)(p_p * H * (1.0 / S));
    k->rho = rho_p + K * e;
    v_rho = (float)((1.0 - K * H) * p_p);
    k->sigma_rho = sqrtf(v_rho);
    if (k->rho < RHO_MIN) {
      k->sigma_rho += RHO_MIN - k->rho;
      k->rho = RHO_MIN;
    } else if (k->rho > RHO_MAX) {
      k->rho = RHO_MAX;
    } else if (isnan(k->rho) || isnan(k->sigma_rho) || isinf(k->rho) || isinf(k->sigma_rho)) {
      k->rho = RHO_INIT;
      k->sigma_rho = RHO_MAX;
    }
  }
}
