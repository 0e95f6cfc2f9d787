/* TEST INFRASTRUCTURE — not part of the product.  See glia_oracle.h.
 *
 * CPU restatement of GLIA's hmt feature + merge-order path.  Each function
 * cites the reference code it follows (paths relative to
 * /root/reference/code).  The restatement works on mergeable statistics
 * (per base region, per directed base edge) instead of pixel lists; that this
 * reproduces the reference exactly is what tests/test_oracle_vs_ref.py pins
 * against oracle/_ref (the unmodified reference headers).
 */
#include "glia_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>

#define FEPS 2.22e-16 /* glia_base.hxx:59 */
static double sdivide(double a, double b, double d) { return fabs(b) >= FEPS ? a / b : d; } /* :77 */
static double slog(double x, double d) { return x > 0.0 ? log(x) : d; }                      /* :80 */
static double ssqrt(double x, double d) { return x >= 0.0 ? sqrt(x) : d; }                   /* :83 */

/* mergeable statistics of a voxel multiset w.r.t. one image */
typedef struct {
  uint64_t n;
  uint64_t* hist; /* bins */
  double s, s2;
  double mn, mx; /* DBL_MAX / -DBL_MAX when n == 0 */
} stat_t;

typedef struct { int bins; double lo, hi; double* bounds; const float* img; } imgspec;

struct go_state {
  go_cfg cfg;
  int D, R, E, C;
  size_t nvox;
  uint32_t* label;          /* [R] ascending */
  uint32_t* cnt; uint32_t* border;
  int32_t* bbmin; int32_t* bbmax; /* [R][3] */
  imgspec rs[GO_MAX_IMG], rls[GO_MAX_IMG], bs[GO_MAX_IMG];
  stat_t* rstat[GO_MAX_IMG];  /* [R] */
  stat_t* rlstat[GO_MAX_IMG]; /* [R] */
  /* directed edges sorted by (a,b) index */
  int* ea; int* eb; uint32_t* ecnt; int* rev; /* rev[e] = index of (b,a) or -1 */
  uint32_t* ethr;           /* [E][T] */
  stat_t* estat[GO_MAX_IMG]; /* [E] */
  int* row;                 /* CSR by a: [R+1] */
  /* incoming edge lists (for non-mutual lookups): in_row [R+1], in_e */
  double normArea, normLen;
  /* GLIA_USE_MEDIAN_AS_FEATS: the values themselves, per base region / per directed base edge */
  float** rvals[GO_MAX_IMG]; /* [R] -> cnt[r] values */
  float** evals[GO_MAX_IMG]; /* [E] -> ecnt[e] values */
  /* scratch */
  int* mark; int stamp;
  int* mark2;
};

/* growing value list of a voxel set (median variant) */
typedef struct { float* v; size_t n, cap; } vlist;
static void vl_add(vlist* l, const float* src, size_t n) {
  if (l->n + n > l->cap) { l->cap = 2 * (l->n + n) + 16; l->v = (float*)realloc(l->v, 4 * l->cap); }
  memcpy(l->v + l->n, src, 4 * n); l->n += n;
}
static int cmp_f32(const void* a, const void* b) {
  float x = *(const float*)a, y = *(const float*)b;
  return x < y ? -1 : x > y;
}

static void stat_init(stat_t* s, int bins) {
  s->n = 0; s->s = s->s2 = 0; s->mn = DBL_MAX; s->mx = -DBL_MAX;
  s->hist = (uint64_t*)calloc((size_t)(bins > 0 ? bins : 1), 8);
}
static void stat_clear(stat_t* s, int bins) {
  s->n = 0; s->s = s->s2 = 0; s->mn = DBL_MAX; s->mx = -DBL_MAX;
  memset(s->hist, 0, (size_t)bins * 8);
}
static void stat_merge(stat_t* d, const stat_t* s, int bins) {
  d->n += s->n; d->s += s->s; d->s2 += s->s2;
  if (s->mn < d->mn) d->mn = s->mn;
  if (s->mx > d->mx) d->mx = s->mx;
  for (int i = 0; i < bins; ++i) d->hist[i] += s->hist[i];
}

/* util/image_stats.hxx:12-36 — cumulative bounds, not offset by lo; a value in
 * (lo,hi) that is >= the last bound falls in no bin. Returns bin or -1. */
static int hist_bin(const imgspec* sp, float val) {
  double v = (double)val;
  if (v > sp->lo && v < sp->hi) {
    for (int i = 0; i < sp->bins; ++i) if (v < sp->bounds[i]) return i;
    return -1;
  } else if (v <= sp->lo) return 0;
  return sp->bins - 1;
}
static void stat_add(stat_t* s, const imgspec* sp, float val) {
  s->n++;
  int b = hist_bin(sp, val);
  if (b >= 0) s->hist[b]++;
  s->s += val;                       /* type/feat.hxx:730 */
  s->s2 += (double)val * val;        /* :731 */
  if (val < s->mn) s->mn = val;
  if (val > s->mx) s->mx = val;
}
static void spec_init(imgspec* sp, const float* img, int bins, double lo, double hi) {
  sp->img = img; sp->bins = bins; sp->lo = lo; sp->hi = hi;
  sp->bounds = (double*)malloc(8 * (size_t)(bins > 0 ? bins : 1));
  double interval = (hi - lo) / bins;
  sp->bounds[0] = interval;
  for (int i = 1; i < bins; ++i) sp->bounds[i] = sp->bounds[i - 1] + interval;
}

static int cmp_u32(const void* a, const void* b) {
  uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
  return x < y ? -1 : x > y;
}
static int cmp_u64(const void* a, const void* b) {
  uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return x < y ? -1 : x > y;
}
static int find_u32(const uint32_t* a, int n, uint32_t v) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) { int m = (lo + hi) >> 1; if (a[m] == v) return m; if (a[m] < v) lo = m + 1; else hi = m - 1; }
  return -1;
}
static int find_u64(const uint64_t* a, int n, uint64_t v) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) { int m = (lo + hi) >> 1; if (a[m] == v) return m; if (a[m] < v) lo = m + 1; else hi = m - 1; }
  return -1;
}

/* type/neighbor.hxx:72-126 + util/struct.hxx:95-108: first differing valid
 * neighbour in order d0-,d0+,d1-,d1+,(d2-,d2+); nvalid counts in-image,
 * unmasked neighbours.  Returns 1 boundary (nb set), 2 border, 0 interior. */
static int classify(const go_state* s, const uint32_t* L, const uint32_t* M, size_t p,
                    const int* xyz, uint32_t* nb) {
  const int* dims = s->cfg.dims;
  size_t stride[3] = {1, (size_t)dims[0], (size_t)dims[0] * dims[1]};
  uint32_t me = L[p];
  int nvalid = 0, found = 0;
  for (int d = 0; d < s->D; ++d) {
    for (int sgn = -1; sgn <= 1; sgn += 2) {
      int c = xyz[d] + sgn;
      if (c < 0 || c >= dims[d]) continue;
      size_t q = sgn < 0 ? p - stride[d] : p + stride[d];
      if (M && M[q] == 0) continue;
      ++nvalid;
      if (!found && L[q] != me) { found = 1; *nb = L[q]; }
    }
  }
  if (found) return 1;
  return nvalid < 2 * s->D ? 2 : 0;
}

go_state* go_build(const go_cfg* cfg, const uint32_t* L, const uint32_t* M, const float* pb) {
  go_state* s = (go_state*)calloc(1, sizeof(*s));
  s->cfg = *cfg;
  s->D = cfg->ndim;
  if (s->D == 2) s->cfg.dims[2] = 1;
  const int* dims = s->cfg.dims;
  size_t N = (size_t)dims[0] * dims[1] * dims[2];
  s->nvox = N;
  /* main_merge_order_bc.cxx:36-39; util/image.hxx:173-192 */
  double diag = 0; for (int d = 0; d < s->D; ++d) diag += (double)dims[d] * dims[d];
  s->normArea = cfg->normalize ? (double)N : 1.0;
  s->normLen = cfg->normalize ? sqrt(diag) : 1.0;
  for (int i = 0; i < cfg->n_rimg; ++i) spec_init(&s->rs[i], cfg->rimg[i], cfg->rbins[i], cfg->rlo[i], cfg->rhi[i]);
  for (int i = 0; i < cfg->n_rlimg; ++i) spec_init(&s->rls[i], cfg->rlimg[i], cfg->rlbins[i], cfg->rllo[i], cfg->rlhi[i]);
  for (int i = 0; i < cfg->n_bimg; ++i) spec_init(&s->bs[i], cfg->bimg[i], cfg->bbins[i], cfg->blo[i], cfg->bhi[i]);
  /* labels present (unmasked voxels only): util/struct.hxx:61-92 */
  uint32_t* tmp = (uint32_t*)malloc(4 * (N ? N : 1));
  size_t nu = 0;
  for (size_t p = 0; p < N; ++p) if (!M || M[p] != 0 || cfg->contour_only) tmp[nu++] = L[p];
  qsort(tmp, nu, 4, cmp_u32);
  int R = 0;
  for (size_t i = 0; i < nu; ++i) if (i == 0 || tmp[i] != tmp[i - 1]) tmp[R++] = tmp[i];
  s->R = R;
  s->label = (uint32_t*)malloc(4 * (size_t)(R ? R : 1));
  memcpy(s->label, tmp, 4 * (size_t)R);
  free(tmp);
  s->cnt = (uint32_t*)calloc((size_t)R + 1, 4);
  s->border = (uint32_t*)calloc((size_t)R + 1, 4);
  s->bbmin = (int32_t*)malloc(12 * ((size_t)R + 1));
  s->bbmax = (int32_t*)malloc(12 * ((size_t)R + 1));
  for (int i = 0; i < 3 * R; ++i) { s->bbmin[i] = INT32_MAX; s->bbmax[i] = INT32_MIN; }
  for (int i = 0; i < cfg->n_rimg; ++i) {
    s->rstat[i] = (stat_t*)malloc(sizeof(stat_t) * ((size_t)R + 1));
    for (int r = 0; r < R; ++r) stat_init(&s->rstat[i][r], s->rs[i].bins);
  }
  for (int i = 0; i < cfg->n_rlimg; ++i) {
    s->rlstat[i] = (stat_t*)malloc(sizeof(stat_t) * ((size_t)R + 1));
    for (int r = 0; r < R; ++r) stat_init(&s->rlstat[i][r], s->rls[i].bins);
  }
  /* pass 1: classify every voxel, region stats, collect directed keys */
  uint8_t* cls = (uint8_t*)malloc(N ? N : 1);
  uint64_t* vkey = (uint64_t*)malloc(8 * (N ? N : 1)); /* per boundary voxel */
  size_t nb = 0;
  size_t p = 0;
  int xyz[3];
  for (xyz[2] = 0; xyz[2] < dims[2]; ++xyz[2])
    for (xyz[1] = 0; xyz[1] < dims[1]; ++xyz[1])
      for (xyz[0] = 0; xyz[0] < dims[0]; ++xyz[0], ++p) {
        cls[p] = 0;
        if (M && M[p] == 0 && !cfg->contour_only) continue;
        int a = find_u32(s->label, R, L[p]);
        s->cnt[a]++;
        for (int d = 0; d < s->D; ++d) { /* alg/geometry.hxx:21-39 */
          if (xyz[d] < s->bbmin[3 * a + d]) s->bbmin[3 * a + d] = xyz[d];
          if (xyz[d] > s->bbmax[3 * a + d]) s->bbmax[3 * a + d] = xyz[d];
        }
        for (int i = 0; i < cfg->n_rimg; ++i) stat_add(&s->rstat[i][a], &s->rs[i], s->rs[i].img[p]);
        for (int i = 0; i < cfg->n_rlimg; ++i) stat_add(&s->rlstat[i][a], &s->rls[i], s->rls[i].img[p]);
        uint32_t nbl = 0;
        int c = classify(s, L, M, p, xyz, &nbl);
        if (c == 2) s->border[a]++;
        if (c == 1) {
          int b = find_u32(s->label, R, nbl);
          vkey[p] = ((uint64_t)a << 32) | (uint32_t)b;
          cls[p] = 1;
          ++nb;
        }
      }
  uint64_t* keys = (uint64_t*)malloc(8 * (nb ? nb : 1));
  size_t k = 0;
  for (p = 0; p < N; ++p) if (cls[p]) keys[k++] = vkey[p];
  qsort(keys, nb, 8, cmp_u64);
  int E = 0;
  for (size_t i = 0; i < nb; ++i) if (i == 0 || keys[i] != keys[i - 1]) keys[E++] = keys[i];
  s->E = E;
  int T = cfg->n_thr;
  s->ea = (int*)malloc(4 * ((size_t)E + 1)); s->eb = (int*)malloc(4 * ((size_t)E + 1));
  s->ecnt = (uint32_t*)calloc((size_t)E + 1, 4);
  s->rev = (int*)malloc(4 * ((size_t)E + 1));
  s->ethr = (uint32_t*)calloc(((size_t)E + 1) * (T ? T : 1), 4);
  s->row = (int*)calloc((size_t)R + 2, 4);
  for (int e = 0; e < E; ++e) {
    s->ea[e] = (int)(keys[e] >> 32); s->eb[e] = (int)(keys[e] & 0xffffffffu);
    s->row[s->ea[e] + 1]++;
  }
  for (int r = 0; r < R; ++r) s->row[r + 1] += s->row[r];
  s->C = 0;
  for (int e = 0; e < E; ++e) {
    s->rev[e] = find_u64(keys, E, ((uint64_t)s->eb[e] << 32) | (uint32_t)s->ea[e]);
    if (s->rev[e] >= 0 && s->ea[e] < s->eb[e]) s->C++;
  }
  for (int i = 0; i < cfg->n_bimg; ++i) {
    s->estat[i] = (stat_t*)malloc(sizeof(stat_t) * ((size_t)E + 1));
    for (int e = 0; e < E; ++e) stat_init(&s->estat[i][e], s->bs[i].bins);
  }
  /* pass 2: per directed edge statistics */
  for (p = 0; p < N; ++p) {
    if (!cls[p]) continue;
    int e = find_u64(keys, E, vkey[p]);
    s->ecnt[e]++;
    for (int t = 0; t < T; ++t) if ((double)pb[p] >= cfg->thr[t]) s->ethr[(size_t)e * T + t]++; /* feat.hxx:498 */
    for (int i = 0; i < cfg->n_bimg; ++i) stat_add(&s->estat[i][e], &s->bs[i], s->bs[i].img[p]);
  }
  if (cfg->use_median) { /* type/feat.hxx:711-724: the values of every point of the set */
    size_t* cur = (size_t*)calloc((size_t)(R > E ? R : E) + 1, sizeof(size_t));
    for (int i = 0; i < cfg->n_rimg; ++i) {
      s->rvals[i] = (float**)calloc((size_t)R + 1, sizeof(float*));
      for (int r = 0; r < R; ++r) s->rvals[i][r] = (float*)malloc(4 * (size_t)(s->cnt[r] ? s->cnt[r] : 1));
      memset(cur, 0, sizeof(size_t) * ((size_t)R + 1));
      for (p = 0; p < N; ++p) {
        if (M && M[p] == 0 && !cfg->contour_only) continue;
        int a = find_u32(s->label, R, L[p]);
        s->rvals[i][a][cur[a]++] = s->rs[i].img[p];
      }
    }
    for (int i = 0; i < cfg->n_bimg; ++i) {
      s->evals[i] = (float**)calloc((size_t)E + 1, sizeof(float*));
      for (int e = 0; e < E; ++e) s->evals[i][e] = (float*)malloc(4 * (size_t)(s->ecnt[e] ? s->ecnt[e] : 1));
      memset(cur, 0, sizeof(size_t) * ((size_t)E + 1));
      for (p = 0; p < N; ++p) {
        if (!cls[p]) continue;
        int e = find_u64(keys, E, vkey[p]);
        s->evals[i][e][cur[e]++] = s->bs[i].img[p];
      }
    }
    free(cur);
  }
  free(keys); free(vkey); free(cls);
  s->mark = (int*)calloc((size_t)R + 1, 4);
  s->mark2 = (int*)calloc((size_t)R + 1, 4);
  s->stamp = 0;
  return s;
}

static void stat_free_arr(stat_t* a, int n) {
  if (!a) return;
  for (int i = 0; i < n; ++i) free(a[i].hist);
  free(a);
}
void go_free(go_state* s) {
  if (!s) return;
  for (int i = 0; i < GO_MAX_IMG; ++i) {
    if (s->rvals[i]) { for (int r = 0; r < s->R; ++r) free(s->rvals[i][r]); free(s->rvals[i]); }
    if (s->evals[i]) { for (int e = 0; e < s->E; ++e) free(s->evals[i][e]); free(s->evals[i]); }
    stat_free_arr(s->rstat[i], s->R); stat_free_arr(s->rlstat[i], s->R); stat_free_arr(s->estat[i], s->E);
    free(s->rs[i].bounds); free(s->rls[i].bounds); free(s->bs[i].bounds);
  }
  free(s->label); free(s->cnt); free(s->border); free(s->bbmin); free(s->bbmax);
  free(s->ea); free(s->eb); free(s->ecnt); free(s->rev); free(s->ethr); free(s->row);
  free(s->mark); free(s->mark2); free(s);
}
int go_num_regions(const go_state* s) { return s->R; }
int go_num_edges(const go_state* s) { return s->E; }
int go_num_candidates(const go_state* s) { return s->C; }

void go_get_regions(const go_state* s, uint32_t* label, uint32_t* count, uint32_t* border,
                    int32_t* bbmin, int32_t* bbmax) {
  if (label) memcpy(label, s->label, 4 * (size_t)s->R);
  if (count) memcpy(count, s->cnt, 4 * (size_t)s->R);
  if (border) memcpy(border, s->border, 4 * (size_t)s->R);
  if (bbmin) memcpy(bbmin, s->bbmin, 12 * (size_t)s->R);
  if (bbmax) memcpy(bbmax, s->bbmax, 12 * (size_t)s->R);
}
static void dump_stats(const stat_t* st, int n, int bins, uint32_t* hist, double* sum, double* sum2,
                       float* mn, float* mx) {
  for (int i = 0; i < n; ++i) {
    if (hist) for (int b = 0; b < bins; ++b) hist[(size_t)i * bins + b] = (uint32_t)st[i].hist[b];
    if (sum) sum[i] = st[i].s;
    if (sum2) sum2[i] = st[i].s2;
    if (mn) mn[i] = (float)st[i].mn;
    if (mx) mx[i] = (float)st[i].mx;
  }
}
void go_get_region_stats(const go_state* s, int kind, int i, uint32_t* hist, double* sum,
                         double* sum2, float* mn, float* mx) {
  if (kind == 0) dump_stats(s->rstat[i], s->R, s->rs[i].bins, hist, sum, sum2, mn, mx);
  else dump_stats(s->rlstat[i], s->R, s->rls[i].bins, hist, sum, sum2, mn, mx);
}
void go_get_edges(const go_state* s, uint32_t* a, uint32_t* b, uint32_t* count, uint8_t* mutual,
                  uint32_t* thr_count) {
  for (int e = 0; e < s->E; ++e) {
    if (a) a[e] = s->label[s->ea[e]];
    if (b) b[e] = s->label[s->eb[e]];
    if (count) count[e] = s->ecnt[e];
    if (mutual) mutual[e] = s->rev[e] >= 0;
  }
  if (thr_count) memcpy(thr_count, s->ethr, 4 * (size_t)s->E * s->cfg.n_thr);
}
void go_get_edge_stats(const go_state* s, int i, uint32_t* hist, double* sum, double* sum2,
                       float* mn, float* mx) {
  dump_stats(s->estat[i], s->E, s->bs[i].bins, hist, sum, sum2, mn, mx);
}

/* ------------------------------------------------------------------------
 * Features of a region given as a set of base regions (type/region.hxx:66-75,
 * type/feat.hxx:71-90, :485-502, :632-638, :706-738; hmt/bc_feat.hxx:88-127).
 * ---------------------------------------------------------------------- */
typedef struct {
  double entropy, mean, stddev, mn, mx, median;
  double hist[256];
} imgfeat;

/* ImageRealFeats::generate under GLIA_USE_MEDIAN_AS_FEATS (type/feat.hxx:711-724): amedian = the n/2-th order
 * statistic (util/stats.hxx:83-91), mean = sum / n in double (:57-59), var = two-pass sum of squared deviations / n
 * (:62-71).  Consumes (sorts) the list. */
static void real_feats_from_values(imgfeat* f, vlist* l) {
  f->median = 0.0;
  if (l->n == 0) return; /* generate() returns before touching anything */
  double sum = 0.0;
  for (size_t i = 0; i < l->n; ++i) sum += l->v[i];
  f->mean = sdivide(sum, (double)l->n, 0.0);
  double var = 0.0;
  for (size_t i = 0; i < l->n; ++i) { double dx = l->v[i] - f->mean; var += dx * dx; }
  f->stddev = ssqrt(sdivide(var, (double)l->n, 0.0), 0.0);
  qsort(l->v, l->n, 4, cmp_f32);
  f->mn = l->v[0]; f->mx = l->v[l->n - 1];
  f->median = l->v[l->n / 2];
}

typedef struct {
  double area, perim, compactness, bboxArea, bboxSize[3];
  double validPerims[GO_MAX_THR], rValidPerims[GO_MAX_THR];
  imgfeat region[GO_MAX_IMG], labelRegion[GO_MAX_IMG], boundary[GO_MAX_IMG];
  int has_sal; double sal;
} regfeat;

/* util/image_stats.hxx:39-53, util/stats.hxx:145-152, type/feat.hxx:706-738 */
static void img_feats(imgfeat* f, const stat_t* st, int bins) {
  uint32_t n = (uint32_t)st->n;
  f->entropy = 0.0;
  for (int i = 0; i < bins; ++i) f->hist[i] = n ? st->hist[i] / (double)n : 0.0;
  for (int i = 0; i < bins; ++i)
    if (!(fabs(f->hist[i] - 0.0) < FEPS)) f->entropy -= f->hist[i] * log2(f->hist[i]);
  f->median = 0.0;
  if (n == 0) { f->mean = f->stddev = f->mn = f->mx = 0.0; return; }
  int ni = (int)n;
  f->mean = st->s / ni;
  f->stddev = ssqrt(st->s2 / ni - f->mean * f->mean, 0.0);
  f->mn = st->mn; f->mx = st->mx;
}

static stat_t g_tmp; static int g_tmp_bins = 0;
static stat_t* tmp_stat(int bins) {
  if (g_tmp_bins < 256) { free(g_tmp.hist); g_tmp.hist = (uint64_t*)calloc(256, 8); g_tmp_bins = 256; }
  stat_clear(&g_tmp, bins > 0 ? bins : 1);
  return &g_tmp;
}

/* is directed edge e of a region (membership marks == stamp) still an entry? */
#define LIVE(s, e, mk, st) (!((mk)[(s)->eb[e]] == (st) && (s)->rev[e] >= 0))

static void region_feats(go_state* s, const int* S, int n, regfeat* f, const double* sal) {
  const go_cfg* c = &s->cfg;
  int D = s->D, T = c->n_thr;
  int st = ++s->stamp;
  for (int i = 0; i < n; ++i) s->mark[S[i]] = st;
  uint64_t area = 0, nbnd = 0, nborder = 0, vp[GO_MAX_THR] = {0};
  int32_t mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  for (int i = 0; i < n; ++i) {
    int a = S[i];
    area += s->cnt[a]; nborder += s->border[a];
    for (int d = 0; d < D; ++d) {
      if (s->bbmin[3 * a + d] < mn[d]) mn[d] = s->bbmin[3 * a + d];
      if (s->bbmax[3 * a + d] > mx[d]) mx[d] = s->bbmax[3 * a + d];
    }
    for (int e = s->row[a]; e < s->row[a + 1]; ++e) {
      if (!LIVE(s, e, s->mark, st)) continue;
      nbnd += s->ecnt[e];
      for (int t = 0; t < T; ++t) vp[t] += s->ethr[(size_t)e * T + t];
    }
  }
  /* RegionShapeFeats::generate, type/feat.hxx:71-90 */
  f->area = (double)area;
  f->perim = (double)(nbnd + nborder);
  f->compactness = sdivide(pow(f->perim, (double)D / (D - 1)), f->area, 0.0);
  f->area = sdivide(f->area, s->normArea, 0.0);
  f->perim = sdivide(f->perim, s->normLen, 0.0);
  f->bboxArea = 1.0;
  for (int d = 0; d < D; ++d) {
    double bb = n ? (double)(mx[d] - mn[d]) : 0.0; /* geometry.hxx:36-38: max - min */
    f->bboxSize[d] = sdivide(bb, s->normLen, 0.0);
    f->bboxArea *= bb;
  }
  f->bboxArea = sdivide(f->bboxArea, s->normArea, 0.0);
  /* ImageRegionShapeFeats::generate, type/feat.hxx:485-502 */
  for (int t = 0; t < T; ++t) {
    f->validPerims[t] = sdivide((double)(uint32_t)vp[t], s->normLen, 0.0);
    f->rValidPerims[t] = sdivide((double)(uint32_t)vp[t], (double)nbnd, 0.0);
  }
  for (int i = 0; i < c->n_rimg; ++i) {
    stat_t* t = tmp_stat(s->rs[i].bins);
    for (int j = 0; j < n; ++j) stat_merge(t, &s->rstat[i][S[j]], s->rs[i].bins);
    img_feats(&f->region[i], t, s->rs[i].bins);
    if (c->use_median) {
      vlist l = {0, 0, 0};
      for (int j = 0; j < n; ++j) vl_add(&l, s->rvals[i][S[j]], s->cnt[S[j]]);
      real_feats_from_values(&f->region[i], &l);
      free(l.v);
    }
  }
  for (int i = 0; i < c->n_rlimg; ++i) {
    stat_t* t = tmp_stat(s->rls[i].bins);
    for (int j = 0; j < n; ++j) stat_merge(t, &s->rlstat[i][S[j]], s->rls[i].bins);
    img_feats(&f->labelRegion[i], t, s->rls[i].bins);
  }
  for (int i = 0; i < c->n_bimg; ++i) {
    stat_t* t = tmp_stat(s->bs[i].bins);
    for (int j = 0; j < n; ++j)
      for (int e = s->row[S[j]]; e < s->row[S[j] + 1]; ++e)
        if (LIVE(s, e, s->mark, st)) stat_merge(t, &s->estat[i][e], s->bs[i].bins);
    img_feats(&f->boundary[i], t, s->bs[i].bins);
    if (c->use_median) {
      vlist l = {0, 0, 0};
      for (int j = 0; j < n; ++j)
        for (int e = s->row[S[j]]; e < s->row[S[j] + 1]; ++e)
          if (LIVE(s, e, s->mark, st)) vl_add(&l, s->evals[i][e], s->ecnt[e]);
      real_feats_from_values(&f->boundary[i], &l);
      free(l.v);
    }
  }
  f->has_sal = sal != NULL;
  f->sal = sal ? *sal : 0.0;
}

/* util/stats.hxx:155-185 (pow(d,2) == d*d) */
static double dist_l1(const double* a, const double* b, int n) {
  double r = 0; for (int i = 0; i < n; ++i) r += fabs(a[i] - b[i]); return r;
}
static double dist_x2(const double* a, const double* b, int n) {
  double r = 0;
  for (int i = 0; i < n; ++i) { double d = a[i] - b[i]; r += d * d / (a[i] + b[i] + FEPS); }
  return r;
}

/* ImageFeats::serialize = ImageLabelFeats ([histogram] entropy, type/feat.hxx:616-623) then ImageRealFeats
 * ([median] mean stddev min max, :684-694); real = 0: ImageLabelFeats alone. */
static double* push_img(const go_cfg* c, const imgfeat* g, int bins, int real, double* p) {
  if (c->use_hist) for (int b = 0; b < bins; ++b) *p++ = g->hist[b];
  *p++ = g->entropy;
  if (!real) return p;
  if (c->use_median) *p++ = g->median;
  *p++ = g->mean; *p++ = g->stddev; *p++ = g->mn; *p++ = g->mx;
  return p;
}

static void push_region(const go_state* s, const regfeat* f, double** o) {
  const go_cfg* c = &s->cfg;
  int lg = c->use_log;
  double* p = *o;
  *p++ = lg ? slog(f->area, 0) : f->area;
  *p++ = lg ? slog(f->perim, 0) : f->perim;
  *p++ = f->compactness;
  *p++ = lg ? slog(f->bboxArea, 0) : f->bboxArea;
  for (int d = 0; d < s->D; ++d) *p++ = lg ? slog(f->bboxSize[d], 0) : f->bboxSize[d];
  for (int t = 0; t < c->n_thr; ++t) *p++ = lg ? slog(f->validPerims[t], 0) : f->validPerims[t];
  for (int t = 0; t < c->n_thr; ++t) *p++ = f->rValidPerims[t];
  for (int i = 0; i < c->n_rimg; ++i) p = push_img(c, &f->region[i], s->rs[i].bins, 1, p);
  for (int i = 0; i < c->n_rlimg; ++i) p = push_img(c, &f->labelRegion[i], s->rls[i].bins, 0, p);
  for (int i = 0; i < c->n_bimg; ++i) p = push_img(c, &f->boundary[i], s->bs[i].bins, 1, p);
  if (f->has_sal) *p++ = f->sal;
  *o = p;
}

int go_row_dim(const go_state* s, int with_sal) {
  const go_cfg* c = &s->cfg;
  int real = c->use_median ? 5 : 4;
  if (c->simple) return 5 + c->n_bimg * (c->use_median ? 2 : 1) + 4 * c->n_rimg + 2 * c->n_rlimg;
  int x0 = 11 + 4 * c->n_thr + (3 + real) * c->n_rimg + 3 * c->n_rlimg + (with_sal ? 2 : 0);
  int xi = s->D + 4 + 2 * c->n_thr + (with_sal ? 1 : 0);
  for (int i = 0; i < c->n_rimg; ++i) xi += (c->use_hist ? s->rs[i].bins : 0) + 1 + real;
  for (int i = 0; i < c->n_rlimg; ++i) xi += (c->use_hist ? s->rls[i].bins : 0) + 1;
  for (int i = 0; i < c->n_bimg; ++i) { int l = (c->use_hist ? s->bs[i].bins : 0) + 1 + real; xi += l; x0 += l; }
  return x0 + 3 * xi;
}

/* Candidate row for regions X, Y (base-index sets) and Z = X u Y
 * (main_merge_order_bc.cxx:54-95; bc_feat.hxx:179-238; feat.hxx:124-132,
 * :176-186, :567-589, :663-669, :801-810; util/struct.hxx:10-16;
 * region.hxx:42-51).  f0,f1,f2 are the RegionFeats of X, Y, Z. */
static void candidate_row(go_state* s, const int* X, int nx, const int* Y, int ny,
                          const regfeat* fX, const regfeat* fY, const regfeat* fZ, double* out) {
  const go_cfg* c = &s->cfg;
  int T = c->n_thr, lg = c->use_log;
  const regfeat *f1 = fX, *f2 = fY;
  if (f1->area > f2->area) { const regfeat* t = f1; f1 = f2; f2 = t; } /* :77-80 */
  /* pair boundary b(X,Y) */
  int sx = ++s->stamp; for (int i = 0; i < nx; ++i) s->mark[X[i]] = sx;
  int sy = ++s->stamp; for (int i = 0; i < ny; ++i) s->mark2[Y[i]] = sy;
  /* active flags: base region with at least one live entry, stored as sign in a scratch pass */
  /* (computed on the fly: small sets) */
  uint64_t nb = 0, vp[GO_MAX_THR] = {0};
  stat_t bst[GO_MAX_IMG];
  vlist bvl[GO_MAX_IMG];
  memset(bvl, 0, sizeof(bvl));
  for (int i = 0; i < c->n_bimg; ++i) stat_init(&bst[i], s->bs[i].bins);
  for (int side = 0; side < 2; ++side) {
    const int* A = side ? Y : X; int na = side ? ny : nx;
    int* mkA = side ? s->mark2 : s->mark; int stA = side ? sy : sx;
    int* mkB = side ? s->mark : s->mark2; int stB = side ? sx : sy;
    for (int i = 0; i < na; ++i) {
      for (int e = s->row[A[i]]; e < s->row[A[i] + 1]; ++e) {
        if (!LIVE(s, e, mkA, stA)) continue;
        int b = s->eb[e];
        if (mkB[b] != stB) continue;
        /* b must be the first key of some remaining entry of the other region */
        int act = 0;
        for (int e2 = s->row[b]; e2 < s->row[b + 1] && !act; ++e2) act = LIVE(s, e2, mkB, stB);
        if (!act) continue;
        nb += s->ecnt[e];
        for (int t = 0; t < T; ++t) vp[t] += s->ethr[(size_t)e * T + t];
        for (int k = 0; k < c->n_bimg; ++k) stat_merge(&bst[k], &s->estat[k][e], s->bs[k].bins);
        if (c->use_median) for (int k = 0; k < c->n_bimg; ++k) vl_add(&bvl[k], s->evals[k][e], s->ecnt[e]);
      }
    }
  }
  double* p = out;
  double areaDiff = fabs(f1->area - f2->area), perimDiff = fabs(f1->perim - f2->perim);
  double bl = sdivide(ceil((uint32_t)nb / 2.0), s->normLen, 0.0);
  imgfeat bf[GO_MAX_IMG];
  for (int k = 0; k < c->n_bimg; ++k) {
    img_feats(&bf[k], &bst[k], s->bs[k].bins); free(bst[k].hist);
    if (c->use_median) { real_feats_from_values(&bf[k], &bvl[k]); free(bvl[k].v); }
  }
  if (c->simple) { /* bc_feat.hxx:247-279 (log applied before selection) */
    *p++ = lg ? slog(f1->area, 0) : f1->area;
    *p++ = lg ? slog(f2->area, 0) : f2->area;
    *p++ = lg ? slog(f1->perim, 0) : f1->perim;
    *p++ = lg ? slog(f2->perim, 0) : f2->perim;
    *p++ = lg ? slog(bl, 0) : bl;
    for (int k = 0; k < c->n_bimg; ++k) { *p++ = bf[k].mean; if (c->use_median) *p++ = bf[k].median; } /* bc_feat.hxx:263-268 */
    for (int i = 0; i < c->n_rimg; ++i) {
      const imgfeat *a = &f1->region[i], *b = &f2->region[i];
      *p++ = fabs(a->mean - b->mean);
      *p++ = dist_l1(a->hist, b->hist, s->rs[i].bins);
      *p++ = dist_x2(a->hist, b->hist, s->rs[i].bins);
      *p++ = fabs(a->entropy - b->entropy);
    }
    for (int i = 0; i < c->n_rlimg; ++i) {
      const imgfeat *a = &f1->labelRegion[i], *b = &f2->labelRegion[i];
      *p++ = dist_l1(a->hist, b->hist, s->rls[i].bins);
      *p++ = dist_x2(a->hist, b->hist, s->rls[i].bins);
    }
    return;
  }
  *p++ = lg ? slog(areaDiff, 0) : areaDiff;
  *p++ = sdivide(areaDiff, f1->area, 0.0);
  *p++ = sdivide(areaDiff, f2->area, 0.0);
  *p++ = lg ? slog(perimDiff, 0) : perimDiff;
  *p++ = sdivide(perimDiff, f1->perim, 0.0);
  *p++ = sdivide(perimDiff, f2->perim, 0.0);
  *p++ = lg ? slog(bl, 0) : bl;
  *p++ = sdivide(bl, f1->area, 0.0);
  *p++ = sdivide(bl, f2->area, 0.0);
  *p++ = sdivide(bl, f1->perim, 0.0);
  *p++ = sdivide(bl, f2->perim, 0.0);
  double vbl[GO_MAX_THR];
  for (int t = 0; t < T; ++t) vbl[t] = sdivide(ceil((uint32_t)vp[t] / 2.0), s->normLen, 0.0);
  for (int t = 0; t < T; ++t) *p++ = lg ? slog(vbl[t], 0) : vbl[t];
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], bl, 0.0);
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], f1->perim, 0.0);
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], f2->perim, 0.0);
  for (int i = 0; i < c->n_rimg; ++i) {
    const imgfeat *a = &f1->region[i], *b = &f2->region[i];
    *p++ = dist_l1(a->hist, b->hist, s->rs[i].bins);
    *p++ = dist_x2(a->hist, b->hist, s->rs[i].bins);
    *p++ = fabs(a->entropy - b->entropy);
    if (c->use_median) *p++ = fabs(a->median - b->median); /* feat.hxx:803-805 */
    *p++ = fabs(a->mean - b->mean);
    *p++ = fabs(a->stddev - b->stddev);
    *p++ = fabs(a->mn - b->mn);
    *p++ = fabs(a->mx - b->mx);
  }
  for (int i = 0; i < c->n_rlimg; ++i) {
    const imgfeat *a = &f1->labelRegion[i], *b = &f2->labelRegion[i];
    *p++ = dist_l1(a->hist, b->hist, s->rls[i].bins);
    *p++ = dist_x2(a->hist, b->hist, s->rls[i].bins);
    *p++ = fabs(a->entropy - b->entropy);
  }
  for (int k = 0; k < c->n_bimg; ++k) p = push_img(c, &bf[k], s->bs[k].bins, 1, p);
  if (f1->has_sal && f2->has_sal && fZ->has_sal) { /* bc_feat.hxx:208-213 */
    double d02 = fabs(f1->sal - fZ->sal), d12 = fabs(f2->sal - fZ->sal);
    *p++ = d02 < d12 ? d02 : d12;
    *p++ = d02 < d12 ? d12 : d02;
  }
  push_region(s, f1, &p);
  push_region(s, f2, &p);
  push_region(s, fZ, &p);
}

/* ------------------------------------------------------------------------
 * Initial candidates: mutual base pairs in ascending (min,max) label order
 * (type/boundary_table.hxx:91-114; util/struct_merge_bc.hxx:18-23).
 * ---------------------------------------------------------------------- */
void go_init_candidates(go_state* s, uint32_t* keys, double* rows, const uint8_t* swap) {
  int dim = go_row_dim(s, 0), k = 0;
  regfeat fa, fb, fz;
  for (int e = 0; e < s->E; ++e) { /* edges are sorted by (a,b): a<b with reverse present */
    int a = s->ea[e], b = s->eb[e];
    if (!(a < b && s->rev[e] >= 0)) continue;
    if (keys) { keys[2 * k] = s->label[a]; keys[2 * k + 1] = s->label[b]; }
    if (rows) {
      int Z[2] = {a, b};
      region_feats(s, &a, 1, &fa, NULL);
      region_feats(s, &b, 1, &fb, NULL);
      region_feats(s, Z, 2, &fz, NULL);
      /* The reference passes the pair in the direction it meets first while
       * iterating an unordered_map (boundary_table.hxx:93-104); that only matters
       * when both areas are equal (x1 = first argument).  swap[k] selects (b,a). */
      if (swap && swap[k]) candidate_row(s, &b, 1, &a, 1, &fb, &fa, &fz, rows + (size_t)k * dim);
      else candidate_row(s, &a, 1, &b, 1, &fa, &fb, &fz, rows + (size_t)k * dim);
    }
    ++k;
  }
}

/* ------------------------------------------------------------------------
 * Greedy merge order (util/struct_merge.hxx:10-33; boundary_table.hxx:47-52,
 * :121-167).  Table = ordered map keyed (r0<r1); priority = (saliency,
 * insertion sequence), max first (multimap rbegin: largest key, and among
 * equal keys the most recently inserted).
 * ---------------------------------------------------------------------- */
typedef struct { uint32_t r0, r1; double sal; uint64_t seq; int alive; double* row; } titem;
typedef struct { titem* it; int n, cap; } ttable; /* kept sorted by (r0,r1) */

static int t_lower(const ttable* t, uint32_t r0, uint32_t r1) {
  int lo = 0, hi = t->n;
  while (lo < hi) {
    int m = (lo + hi) >> 1;
    const titem* x = &t->it[m];
    if (x->r0 < r0 || (x->r0 == r0 && x->r1 < r1)) lo = m + 1; else hi = m;
  }
  return lo;
}
static int t_find(const ttable* t, uint32_t r0, uint32_t r1) {
  int i = t_lower(t, r0, r1);
  return (i < t->n && t->it[i].r0 == r0 && t->it[i].r1 == r1) ? i : -1;
}
static int t_insert(ttable* t, titem v) {
  if (t->n == t->cap) { t->cap = t->cap ? 2 * t->cap : 64; t->it = (titem*)realloc(t->it, sizeof(titem) * t->cap); }
  int i = t_lower(t, v.r0, v.r1);
  memmove(&t->it[i + 1], &t->it[i], sizeof(titem) * (size_t)(t->n - i));
  t->it[i] = v; t->n++;
  return i;
}
static void t_erase(ttable* t, int i) {
  free(t->it[i].row);
  memmove(&t->it[i], &t->it[i + 1], sizeof(titem) * (size_t)(t->n - i - 1));
  t->n--;
}

typedef struct { int* base; int n; } rset; /* base-region index set of a (merged) region */

int go_greedy(go_state* s, const rf_oracle* rf, go_pred_fn fn, void* ctx, uint32_t* order,
              double* sal, double* rows, uint32_t* cand_keys, double* cand_rows,
              double* cand_scores, size_t max_cand, size_t* n_cand_out, const uint8_t* init_swap) {
  int dim = go_row_dim(s, 0), R = s->R;
  if (R == 0) { if (n_cand_out) *n_cand_out = 0; return 0; }
  uint32_t maxKey = s->label[R - 1];
  /* region id -> set; ids: base labels then maxKey+1+i */
  int nsets = 2 * R;
  rset* sets = (rset*)calloc((size_t)nsets, sizeof(rset));
  for (int r = 0; r < R; ++r) { sets[r].base = (int*)malloc(4); sets[r].base[0] = r; sets[r].n = 1; }
#define SET_OF(key) ((key) <= maxKey ? &sets[find_u32(s->label, R, (key))] : &sets[R + ((key) - maxKey - 1)])
  ttable tb = {0, 0, 0};
  uint64_t seq = 0; size_t ncand = 0;
  double* row = (double*)malloc(8 * (size_t)dim);
  regfeat fa, fb, fz;
#define FEATURIZE(k0, k1, k2key)                                                        \
  do {                                                                                  \
    rset* A_ = SET_OF(k0); rset* B_ = SET_OF(k1);                                       \
    int* Z_ = (int*)malloc(4 * (size_t)(A_->n + B_->n));                                \
    memcpy(Z_, A_->base, 4 * (size_t)A_->n); memcpy(Z_ + A_->n, B_->base, 4 * (size_t)B_->n); \
    region_feats(s, A_->base, A_->n, &fa, NULL);                                        \
    region_feats(s, B_->base, B_->n, &fb, NULL);                                        \
    region_feats(s, Z_, A_->n + B_->n, &fz, NULL);                                      \
    candidate_row(s, A_->base, A_->n, B_->base, B_->n, &fa, &fb, &fz, row);             \
    free(Z_);                                                                           \
    if (ncand < max_cand) {                                                             \
      if (cand_keys) { cand_keys[3 * ncand] = (k0); cand_keys[3 * ncand + 1] = (k1); cand_keys[3 * ncand + 2] = (k2key); } \
      if (cand_rows) memcpy(cand_rows + ncand * dim, row, 8 * (size_t)dim);             \
    }                                                                                   \
  } while (0)
#define SCORE() (rf ? rf_oracle_predict1(rf, row, dim, -1) : fn(row, dim, ctx))
  /* init: featurise in rmap iteration order (unspecified) — we use key order; scores in key order */
  for (int e = 0; e < s->E; ++e) {
    int a = s->ea[e], b = s->eb[e];
    if (!(a < b && s->rev[e] >= 0)) continue;
    titem v; v.r0 = s->label[a]; v.r1 = s->label[b]; v.alive = 1;
    if (init_swap && init_swap[ncand]) FEATURIZE(v.r1, v.r0, 0u);
    else FEATURIZE(v.r0, v.r1, 0u);
    v.row = (double*)malloc(8 * (size_t)dim); memcpy(v.row, row, 8 * (size_t)dim);
    v.sal = SCORE(); v.seq = seq++;
    if (ncand < max_cand && cand_scores) cand_scores[ncand] = v.sal;
    ++ncand;
    t_insert(&tb, v);
  }
  int M = 0;
  uint32_t keyToAssign = maxKey + 1;
  while (tb.n > 0) {
    int top = 0;
    for (int i = 1; i < tb.n; ++i)
      if (tb.it[i].sal > tb.it[top].sal || (tb.it[i].sal == tb.it[top].sal && tb.it[i].seq > tb.it[top].seq)) top = i;
    uint32_t r0 = tb.it[top].r0, r1 = tb.it[top].r1, r2 = keyToAssign++;
    order[3 * M] = r0; order[3 * M + 1] = r1; order[3 * M + 2] = r2;
    sal[M] = tb.it[top].sal;
    if (rows) memcpy(rows + (size_t)M * dim, tb.it[top].row, 8 * (size_t)dim);
    ++M;
    { /* rmap.merge(r0, r1, r2): region_map.hxx:113-118 */
      rset* A = SET_OF(r0); rset* B = SET_OF(r1); rset* Z = &sets[R + (r2 - maxKey - 1)];
      Z->n = A->n + B->n; Z->base = (int*)malloc(4 * (size_t)Z->n);
      memcpy(Z->base, A->base, 4 * (size_t)A->n); memcpy(Z->base + A->n, B->base, 4 * (size_t)B->n);
    }
    t_erase(&tb, top);
    /* update: scan keys ascending while first <= r1 (boundary_table.hxx:127-166) */
    int i = 0;
    while (i < tb.n && tb.it[i].r0 <= r1) {
      uint32_t k0 = tb.it[i].r0, k1 = tb.it[i].r1, rs;
      if (k0 == r0 || k0 == r1) rs = k1;
      else if (k1 == r0 || k1 == r1) rs = k0;
      else { ++i; continue; }
      titem v; v.r0 = rs; v.r1 = r2; v.alive = 1;
      FEATURIZE(rs, r2, 0u);
      v.row = (double*)malloc(8 * (size_t)dim); memcpy(v.row, row, 8 * (size_t)dim);
      v.sal = SCORE(); v.seq = seq++;
      if (ncand < max_cand && cand_scores) cand_scores[ncand] = v.sal;
      ++ncand;
      t_insert(&tb, v); /* (rs, r2) sorts after every key with first == rs and second < r2 */
      /* erase (rs,r0) and (rs,r1) whichever exist; continue from the element after the current */
      uint32_t ck0 = k0, ck1 = k1;
      int j;
      j = t_find(&tb, rs < r0 ? rs : r0, rs < r0 ? r0 : rs); if (j >= 0) t_erase(&tb, j);
      j = t_find(&tb, rs < r1 ? rs : r1, rs < r1 ? r1 : rs); if (j >= 0) t_erase(&tb, j);
      i = t_lower(&tb, ck0, ck1); /* first key >= the erased current key */
    }
  }
  for (int r = 0; r < nsets; ++r) free(sets[r].base);
  free(sets); free(row); free(tb.it);
  if (n_cand_out) *n_cand_out = ncand;
  return M;
#undef SET_OF
#undef FEATURIZE
#undef SCORE
}

/* ------------------------------------------------------------------------
 * bc_feat stage for a given order (hmt/main_bc_feat.cxx:27-112;
 * bc_feat.hxx:12-26 genSaliencyMap).  Region keys in `order` are labels.
 * ---------------------------------------------------------------------- */
int go_bc_feat(go_state* s, const uint32_t* order, int n, const double* sal_in, double init_sal,
               double sal_bias, double* rows) {
  int R = s->R, with_sal = sal_in != NULL;
  int dim = go_row_dim(s, with_sal);
  /* key -> set; merged keys are arbitrary labels given by the file: map via sorted table */
  uint32_t* mkeys = (uint32_t*)malloc(4 * (size_t)(n ? n : 1));
  rset* msets = (rset*)calloc((size_t)(n ? n : 1), sizeof(rset));
  double* msal = (double*)malloc(8 * (size_t)(n ? n : 1));
  int status = 0;
  for (int i = 0; i < n && !status; ++i) {
    rset in[2]; double insal[2];
    for (int j = 0; j < 2; ++j) {
      uint32_t key = order[3 * i + j];
      int found = 0;
      for (int m = i - 1; m >= 0 && !found; --m)
        if (mkeys[m] == key) { in[j] = msets[m]; insal[j] = msal[m]; found = 1; }
      if (!found) {
        int b = find_u32(s->label, R, key);
        if (b < 0) { status = -1; break; }
        static int single[2]; single[j] = b;
        in[j].base = &single[j]; in[j].n = 1; insal[j] = init_sal;
      }
    }
    if (status) break;
    mkeys[i] = order[3 * i + 2];
    msets[i].n = in[0].n + in[1].n;
    msets[i].base = (int*)malloc(4 * (size_t)msets[i].n);
    memcpy(msets[i].base, in[0].base, 4 * (size_t)in[0].n);
    memcpy(msets[i].base + in[0].n, in[1].base, 4 * (size_t)in[1].n);
    msal[i] = with_sal ? sal_in[i] + sal_bias : 0.0;
    regfeat fa, fb, fz;
    region_feats(s, in[0].base, in[0].n, &fa, with_sal ? &insal[0] : NULL);
    region_feats(s, in[1].base, in[1].n, &fb, with_sal ? &insal[1] : NULL);
    region_feats(s, msets[i].base, msets[i].n, &fz, with_sal ? &msal[i] : NULL);
    candidate_row(s, in[0].base, in[0].n, in[1].base, in[1].n, &fa, &fb, &fz, rows + (size_t)i * dim);
  }
  for (int i = 0; i < n; ++i) free(msets[i].base);
  free(mkeys); free(msets); free(msal);
  return status;
}

/* ------------------------------------------------------------------------
 * merge order from the boundary probability alone
 * (hmt/main_merge_order_pb.cxx:14-40; util/struct_merge.hxx:36-136) and
 * pre_merge (gadget/main_pre_merge.cxx:7-87).  Needs bimg[0] == pb (pair
 * boundary sums / values) and, for pre_merge, rimg[0] == pb (region means).
 * Item data = (mean, n) of the pair boundary (struct_merge.hxx:44-56), or the
 * values themselves for the median variant (:100-112); a merge combines the
 * data of the two items it replaces without touching the image again
 * (:62-76, :119-126).  Table / queue semantics as in go_greedy.
 * ---------------------------------------------------------------------- */
typedef struct { uint32_t r0, r1; double sal; uint64_t seq; double mean; long n; float* v; size_t nv; } pitem;
typedef struct { pitem* it; int n, cap; } ptable;
static int p_lower(const ptable* t, uint32_t r0, uint32_t r1) {
  int lo = 0, hi = t->n;
  while (lo < hi) {
    int m = (lo + hi) >> 1;
    const pitem* x = &t->it[m];
    if (x->r0 < r0 || (x->r0 == r0 && x->r1 < r1)) lo = m + 1; else hi = m;
  }
  return lo;
}
static int p_find(const ptable* t, uint32_t r0, uint32_t r1) {
  int i = p_lower(t, r0, r1);
  return (i < t->n && t->it[i].r0 == r0 && t->it[i].r1 == r1) ? i : -1;
}
static void p_insert(ptable* t, pitem v) {
  if (t->n == t->cap) { t->cap = t->cap ? 2 * t->cap : 64; t->it = (pitem*)realloc(t->it, sizeof(pitem) * t->cap); }
  int i = p_lower(t, v.r0, v.r1);
  memmove(&t->it[i + 1], &t->it[i], sizeof(pitem) * (size_t)(t->n - i));
  t->it[i] = v; t->n++;
}
static void p_erase(ptable* t, int i) {
  free(t->it[i].v);
  memmove(&t->it[i], &t->it[i + 1], sizeof(pitem) * (size_t)(t->n - i - 1));
  t->n--;
}
static double values_amedian(const float* v, size_t n) { /* util/stats.hxx:83-91 (n > 0) */
  float* c = (float*)malloc(4 * n);
  memcpy(c, v, 4 * n);
  qsort(c, n, 4, cmp_f32);
  double r = c[n / 2];
  free(c);
  return r;
}
typedef struct {
  const go_state* s; uint32_t maxKey; int nthr; const int* thr; double rpbThr;
  uint64_t* size; double* sum; /* per region id (base index, then R + merge index) */
  double* rpb; uint8_t* has_rpb;
} premerge_ctx;
static int key_id(const go_state* s, uint32_t maxKey, uint32_t key) {
  return key <= maxKey ? find_u32(s->label, s->R, key) : s->R + (int)(key - maxKey - 1);
}
/* gadget/main_pre_merge.cxx:26-76 */
static int premerge_cond(premerge_ctx* c, uint32_t k0, uint32_t k1) {
  int id[2] = {key_id(c->s, c->maxKey, k0), key_id(c->s, c->maxKey, k1)};
  if (c->size[id[0]] > c->size[id[1]]) { int t = id[0]; id[0] = id[1]; id[1] = t; }
  if ((long long)c->size[id[0]] < c->thr[0]) return 1;
  if (c->nthr > 1) {
    for (int k = 0; k < 2; ++k) {
      if ((long long)c->size[id[k]] < c->thr[1]) {
        if (!c->has_rpb[id[k]]) { c->rpb[id[k]] = sdivide(c->sum[id[k]], (double)c->size[id[k]], 0.0); c->has_rpb[id[k]] = 1; }
        if (c->rpb[id[k]] > c->rpbThr) return 1;
      }
    }
  }
  return 0;
}
typedef struct { double sal; uint64_t seq; int idx; } prio;
static int cmp_prio_desc(const void* a, const void* b) {
  const prio* x = (const prio*)a; const prio* y = (const prio*)b;
  if (x->sal != y->sal) return x->sal > y->sal ? -1 : 1;
  return x->seq > y->seq ? -1 : (x->seq < y->seq);
}

static int pb_greedy(go_state* s, int type, premerge_ctx* pc, uint32_t* order, double* sal) {
  int R = s->R;
  if (R == 0 || s->cfg.n_bimg < 1) return 0;
  if (type == 1 && !s->cfg.use_median) return -1;
  uint32_t maxKey = s->label[R - 1];
  ptable tb = {0, 0, 0};
  uint64_t seq = 0;
  /* init: mutual base pairs; pair boundary = (a->b) then (b->a) (util/struct.hxx:10-16) */
  for (int e = 0; e < s->E; ++e) {
    int a = s->ea[e], b = s->eb[e], rv = s->rev[e];
    if (!(a < b && rv >= 0)) continue;
    pitem v; memset(&v, 0, sizeof(v));
    v.r0 = s->label[a]; v.r1 = s->label[b];
    v.n = (long)s->ecnt[e] + (long)s->ecnt[rv];
    v.mean = sdivide(s->estat[0][e].s + s->estat[0][rv].s, (double)v.n, 0.0);
    if (type == 1) {
      v.nv = (size_t)v.n; v.v = (float*)malloc(4 * (v.nv ? v.nv : 1));
      memcpy(v.v, s->evals[0][e], 4 * (size_t)s->ecnt[e]);
      memcpy(v.v + s->ecnt[e], s->evals[0][rv], 4 * (size_t)s->ecnt[rv]);
      v.sal = -values_amedian(v.v, v.nv);
    } else v.sal = -v.mean;
    p_insert(&tb, v);
  }
  for (int i = 0; i < tb.n; ++i) tb.it[i].seq = seq++; /* queue insertion in table order (boundary_table.hxx:108-112) */
  int M = 0;
  uint32_t keyToAssign = maxKey + 1;
  while (tb.n > 0) {
    int top = -1;
    if (!pc) {
      top = 0;
      for (int i = 1; i < tb.n; ++i)
        if (tb.it[i].sal > tb.it[top].sal || (tb.it[i].sal == tb.it[top].sal && tb.it[i].seq > tb.it[top].seq)) top = i;
    } else { /* first item in queue order that passes the condition (boundary_table.hxx:47-52) */
      prio* q = (prio*)malloc(sizeof(prio) * (size_t)tb.n);
      for (int i = 0; i < tb.n; ++i) { q[i].sal = tb.it[i].sal; q[i].seq = tb.it[i].seq; q[i].idx = i; }
      qsort(q, (size_t)tb.n, sizeof(prio), cmp_prio_desc);
      for (int i = 0; i < tb.n && top < 0; ++i) if (premerge_cond(pc, tb.it[q[i].idx].r0, tb.it[q[i].idx].r1)) top = q[i].idx;
      free(q);
      if (top < 0) break;
    }
    uint32_t r0 = tb.it[top].r0, r1 = tb.it[top].r1, r2 = keyToAssign++;
    order[3 * M] = r0; order[3 * M + 1] = r1; order[3 * M + 2] = r2;
    sal[M] = tb.it[top].sal;
    ++M;
    if (pc) { /* rmap.merge(r0, r1, r2): sizes and sums of the merged region */
      int i0 = key_id(s, maxKey, r0), i1 = key_id(s, maxKey, r1), i2 = R + (int)(r2 - maxKey - 1);
      pc->size[i2] = pc->size[i0] + pc->size[i1];
      pc->sum[i2] = pc->sum[i0] + pc->sum[i1];
    }
    p_erase(&tb, top);
    int i = 0;
    while (i < tb.n && tb.it[i].r0 <= r1) { /* boundary_table.hxx:127-166 */
      uint32_t k0 = tb.it[i].r0, k1 = tb.it[i].r1, rs;
      if (k0 == r0 || k0 == r1) rs = k1;
      else if (k1 == r0 || k1 == r1) rs = k0;
      else { ++i; continue; }
      int j0 = p_find(&tb, rs < r0 ? rs : r0, rs < r0 ? r0 : rs);
      int j1 = p_find(&tb, rs < r1 ? rs : r1, rs < r1 ? r1 : rs);
      pitem v; memset(&v, 0, sizeof(v));
      v.r0 = rs; v.r1 = r2;
      double acc = 0.0; long n = 0; /* struct_merge.hxx:62-76: 0s then 1s */
      if (j0 >= 0) { acc += tb.it[j0].mean * tb.it[j0].n; n += tb.it[j0].n; }
      if (j1 >= 0) { acc += tb.it[j1].mean * tb.it[j1].n; n += tb.it[j1].n; }
      v.n = n; v.mean = sdivide(acc, (double)n, 0.0);
      if (type == 1) {
        v.nv = (j0 >= 0 ? tb.it[j0].nv : 0) + (j1 >= 0 ? tb.it[j1].nv : 0);
        v.v = (float*)malloc(4 * (v.nv ? v.nv : 1));
        size_t o = 0;
        if (j0 >= 0) { memcpy(v.v, tb.it[j0].v, 4 * tb.it[j0].nv); o = tb.it[j0].nv; }
        if (j1 >= 0) memcpy(v.v + o, tb.it[j1].v, 4 * tb.it[j1].nv);
        v.sal = -values_amedian(v.v, v.nv);
      } else v.sal = -v.mean;
      v.seq = seq++;
      p_insert(&tb, v);
      int j;
      j = p_find(&tb, rs < r0 ? rs : r0, rs < r0 ? r0 : rs); if (j >= 0) p_erase(&tb, j);
      j = p_find(&tb, rs < r1 ? rs : r1, rs < r1 ? r1 : rs); if (j >= 0) p_erase(&tb, j);
      i = p_lower(&tb, k0, k1);
    }
  }
  for (int i = 0; i < tb.n; ++i) free(tb.it[i].v);
  free(tb.it);
  return M;
}

int go_merge_order_pb(go_state* s, int type, uint32_t* order, double* sal) {
  return pb_greedy(s, type, NULL, order, sal);
}

/* labels_out: the input labels with every merged base region renamed to its final key
 * (transformKeys, util/struct_merge.hxx:188-209; transformImage, util/image.hxx:246-259). */
int go_pre_merge(go_state* s, const int* size_thr, int n_thr, double rpb_thr, uint32_t* order, double* sal,
                 const uint32_t* labels_in, const uint32_t* mask, uint32_t* labels_out) {
  int R = s->R;
  if (R <= 0) return 0;
  if (n_thr < 1 || s->cfg.n_rimg < 1) return -1;
  premerge_ctx c;
  c.s = s; c.maxKey = s->label[R - 1]; c.nthr = n_thr; c.thr = size_thr; c.rpbThr = rpb_thr;
  c.size = (uint64_t*)calloc(2 * (size_t)R, 8); c.sum = (double*)calloc(2 * (size_t)R, 8);
  c.rpb = (double*)calloc(2 * (size_t)R, 8); c.has_rpb = (uint8_t*)calloc(2 * (size_t)R, 1);
  for (int r = 0; r < R; ++r) { c.size[r] = s->cnt[r]; c.sum[r] = s->rstat[0][r].s; }
  int M = pb_greedy(s, 2, &c, order, sal);
  if (labels_out && M >= 0) {
    /* final key of every id: follow the merges */
    uint32_t* fin = (uint32_t*)malloc(4 * 2 * (size_t)R);
    for (int r = 0; r < R; ++r) fin[r] = s->label[r];
    for (int m = 0; m < M; ++m) fin[R + m] = order[3 * m + 2];
    int* up = (int*)malloc(4 * 2 * (size_t)R);
    for (int i = 0; i < 2 * R; ++i) up[i] = i;
    for (int m = 0; m < M; ++m) {
      up[key_id(s, c.maxKey, order[3 * m])] = R + m;
      up[key_id(s, c.maxKey, order[3 * m + 1])] = R + m;
    }
    for (size_t p = 0; p < s->nvox; ++p) {
      labels_out[p] = labels_in[p];
      if (mask && mask[p] == 0) continue;
      int a = find_u32(s->label, R, labels_in[p]);
      if (a < 0) continue;
      int v = a;
      while (up[v] != v) v = up[v];
      labels_out[p] = fin[v];
    }
    free(fin); free(up);
  }
  free(c.size); free(c.sum); free(c.rpb); free(c.has_rpb);
  return M;
}

/* ------------------------------------------------------------------------
 * section-classifier features of region pairs across two 2D
 * sections (hmt/main_sc_feat.cxx:27-117; hmt/sc_feat.hxx:10-170;
 * type/feat.hxx:124-132 shape differences, :191-268 advanced shape (central /
 * Hu moments, eccentricity), :290-370 location, :373-440 set differences;
 * alg/geometry.hxx:44-108; util/image_stats.hxx:308-335 overlaps).
 * Row = RegionPairFeats | RegionFeatsWithLocation(smaller) | ...(larger).
 * ---------------------------------------------------------------------- */
typedef struct { double centroid[2]; double cm[7], hu[7], ecc; } advshape;

/* type/feat.hxx:336-344 (centroid / (size * normLen)), alg/geometry.hxx:44-108, type/feat.hxx:223-243 */
static void adv_shape(const uint32_t* L, const uint32_t* M, const int* dims, uint32_t label, double normLen, advshape* a) {
  double sx = 0, sy = 0; size_t n = 0;
  for (int y = 0; y < dims[1]; ++y)
    for (int x = 0; x < dims[0]; ++x) {
      size_t p = (size_t)y * dims[0] + x;
      if (L[p] != label || (M && M[p] == 0)) continue;
      sx += x; sy += y; ++n;
    }
  double siz = (double)n * normLen;
  a->centroid[0] = sx / siz; a->centroid[1] = sy / siz;
  double ms[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int y = 0; y < dims[1]; ++y)
    for (int x = 0; x < dims[0]; ++x) {
      size_t p = (size_t)y * dims[0] + x;
      if (L[p] != label || (M && M[p] == 0)) continue;
      double dx = x - a->centroid[0], dy = y - a->centroid[1]; /* about the NORMALISED centroid, as the reference does */
      double dx2 = dx * dx, dy2 = dy * dy;
      ms[0] += dy2; ms[1] += dy2 * dy; ms[2] += dx * dy; ms[3] += dx * dy2; ms[4] += dx2; ms[5] += dx2 * dy; ms[6] += dx2 * dx;
    }
  double m00 = (double)n, m002 = m00 * m00, m003 = pow(m00, 2.5), s[7];
  s[0] = sdivide(ms[0], m002, 0.0); s[1] = sdivide(ms[1], m003, 0.0); s[2] = sdivide(ms[2], m002, 0.0);
  s[3] = sdivide(ms[3], m003, 0.0); s[4] = sdivide(ms[4], m002, 0.0); s[5] = sdivide(ms[5], m003, 0.0);
  s[6] = sdivide(ms[6], m003, 0.0);
  if (normLen > 0.0) { /* feat.hxx:231-240: m21 is divided by L^2 there */
    double l2 = normLen * normLen, l3 = l2 * normLen;
    ms[0] /= l2; ms[1] /= l3; ms[2] /= l2; ms[3] /= l3; ms[4] /= l2; ms[5] /= l2; ms[6] /= l3;
  }
  memcpy(a->cm, ms, sizeof(ms));
  double m02 = s[0], m03 = s[1], m11 = s[2], m12 = s[3], m20 = s[4], m21 = s[5], m30 = s[6];
  a->hu[0] = m20 + m02;
  a->hu[1] = pow(m20 - m02, 2) + 4.0 * m11 * m11;
  a->hu[2] = pow(m30 - 3.0 * m12, 2) + pow(3.0 * m21 - m03, 2);
  a->hu[3] = pow(m30 + m12, 2) + pow(m21 + m03, 2);
  a->hu[4] = (m30 - 3.0 * m12) * (m30 + m12) * (pow(m30 + m12, 2) - 3.0 * pow(m21 + m03, 2)) +
             (3.0 * m21 - m03) * (m21 + m03) * (3.0 * pow(m30 + m12, 2) - pow(m21 + m03, 2));
  a->hu[5] = (m20 - m02) * (pow(m30 + m12, 2) - pow(m21 + m03, 2)) + 4.0 * m11 * (m30 + m12) * (m03 + m21);
  a->hu[6] = (3.0 * m21 - m03) * (m12 + m30) * (pow(m30 + m12, 2) - 3.0 * pow(m21 + m03, 2)) -
             (m30 - 3.0 * m12) * (m12 + m03) * (3.0 * pow(m30 + m12, 2) - pow(m21 + m03, 2));
  double ea = a->cm[4] + a->cm[0], eb = ssqrt(pow(a->cm[4] - a->cm[0], 2) + 4.0 * a->cm[2] * a->cm[2], 0.0);
  a->ecc = sdivide(ea + eb, ea - eb, 0.0);
}

int go_sc_row_dim(const go_state* s) {
  const go_cfg* c = &s->cfg;
  int real = c->use_median ? 5 : 4;
  int xi = s->D + 4 + 2 * c->n_thr + 15;
  for (int i = 0; i < c->n_rimg; ++i) xi += (c->use_hist ? s->rs[i].bins : 0) + 1 + real;
  for (int i = 0; i < c->n_rlimg; ++i) xi += (c->use_hist ? s->rls[i].bins : 0) + 1;
  for (int i = 0; i < c->n_bimg; ++i) xi += (c->use_hist ? s->bs[i].bins : 0) + 1 + real;
  return 6 + 1 + 8 + 15 + (3 + real) * (c->n_rimg + c->n_rlimg) + 2 * xi;
}

/* pairs [n][2] = (label in section 0, label in section 1); rows [n][go_sc_row_dim].  s0 / s1 were built from the two
 * sections with the same feature configuration (use_log is taken from the argument, normalisation from norm_*;
 * norm_* <= 0 = image volume / diagonal).  Returns 0, or -1 for an unknown label / unsupported configuration. */
int go_sc_feat(go_state* s0, go_state* s1, const uint32_t* L0, const uint32_t* M0, const uint32_t* L1, const uint32_t* M1,
               const uint32_t* pairs, int n, double norm_area, double norm_len, int use_log, double* rows) {
  if (s0->D != 2 || s1->D != 2 || s0->cfg.use_median || s0->cfg.n_rlimg > s0->cfg.n_rimg) return -1;
  const int* dims = s0->cfg.dims;
  const go_cfg* c = &s0->cfg;
  double vol = (double)dims[0] * dims[1], diag = sqrt((double)dims[0] * dims[0] + (double)dims[1] * dims[1]);
  if (norm_area <= 0.0) norm_area = vol;
  if (norm_len <= 0.0) norm_len = diag;
  go_state* ss[2] = {s0, s1};
  for (int k = 0; k < 2; ++k) { ss[k]->normArea = norm_area; ss[k]->normLen = norm_len; ss[k]->cfg.use_log = use_log; }
  int dim = go_sc_row_dim(s0);
  size_t N = (size_t)dims[0] * dims[1];
  for (int i = 0; i < n; ++i) {
    int a = find_u32(s0->label, s0->R, pairs[2 * i]), b = find_u32(s1->label, s1->R, pairs[2 * i + 1]);
    if (a < 0 || b < 0) return -1;
    regfeat f[2]; advshape as[2];
    region_feats(s0, &a, 1, &f[0], NULL);
    region_feats(s1, &b, 1, &f[1], NULL);
    adv_shape(L0, M0, dims, pairs[2 * i], norm_len, &as[0]);
    adv_shape(L1, M1, dims, pairs[2 * i + 1], norm_len, &as[1]);
    long ov = 0; /* util/image_stats.hxx:308-335: both unmasked, both non-background */
    if (pairs[2 * i] != 0 && pairs[2 * i + 1] != 0)
      for (size_t p = 0; p < N; ++p)
        if (L0[p] == pairs[2 * i] && L1[p] == pairs[2 * i + 1] && (!M0 || M0[p]) && (!M1 || M1[p])) ++ov;
    double* p = rows + (size_t)i * dim;
    /* RegionPairFeats (sc_feat.hxx:119-147), logs as RegionPairFeats::log */
    double areaDiff = fabs(f[0].area - f[1].area), perimDiff = fabs(f[0].perim - f[1].perim);
    *p++ = use_log ? slog(areaDiff, 0) : areaDiff;
    *p++ = sdivide(areaDiff, f[0].area, 0.0); *p++ = sdivide(areaDiff, f[1].area, 0.0);
    *p++ = use_log ? slog(perimDiff, 0) : perimDiff;
    *p++ = sdivide(perimDiff, f[0].perim, 0.0); *p++ = sdivide(perimDiff, f[1].perim, 0.0);
    double d0 = as[0].centroid[0] - as[1].centroid[0], d1 = as[0].centroid[1] - as[1].centroid[1];
    double dist = ssqrt(d0 * d0 + d1 * d1, 0.0);
    if (use_log) { double l = slog(dist, 0.0); dist = l > 0.0 ? l : 0.0; }
    *p++ = dist;
    double o = (double)ov, sd0 = f[0].area - o, sd1 = f[1].area - o, sym = sd0 + sd1; /* areas are the normalised ones */
    *p++ = use_log ? slog(o, 0) : o; *p++ = use_log ? slog(sd0, 0) : sd0; *p++ = use_log ? slog(sd1, 0) : sd1;
    *p++ = use_log ? slog(sym, 0) : sym;
    *p++ = o / f[0].area; *p++ = o / f[1].area; *p++ = sd0 / f[0].area; *p++ = sd1 / f[1].area;
    for (int k = 0; k < 7; ++k) { double d = fabs(as[0].cm[k] - as[1].cm[k]); *p++ = use_log ? slog(d, 0) : d; }
    for (int k = 0; k < 7; ++k) *p++ = fabs(as[0].hu[k] - as[1].hu[k]);
    *p++ = fabs(as[0].ecc - as[1].ecc);
    for (int pass = 0; pass < 2; ++pass) { /* imageRegionDiff, then labelImageRegionDiff - which reads region[i] too (:136-146) */
      int cnt = pass ? c->n_rlimg : c->n_rimg;
      for (int k = 0; k < cnt; ++k) {
        const imgfeat *x = &f[0].region[k], *y = &f[1].region[k];
        *p++ = dist_l1(x->hist, y->hist, s0->rs[k].bins);
        *p++ = dist_x2(x->hist, y->hist, s0->rs[k].bins);
        *p++ = fabs(x->entropy - y->entropy);
        *p++ = fabs(x->mean - y->mean); *p++ = fabs(x->stddev - y->stddev);
        *p++ = fabs(x->mn - y->mn); *p++ = fabs(x->mx - y->mx);
      }
    }
    int first = f[0].area > f[1].area ? 1 : 0; /* keep region 0 area <= region 1 area (:98-101) */
    for (int k = 0; k < 2; ++k) {
      int w = k ? 1 - first : first;
      push_region(ss[w], &f[w], &p);
      for (int j = 0; j < 7; ++j) *p++ = as[w].cm[j];
      for (int j = 0; j < 7; ++j) *p++ = as[w].hu[j];
      *p++ = as[w].ecc;
    }
  }
  return 0;
}
