/* TEST INFRASTRUCTURE — not part of the product.
 *
 * CPU restatement of GLIA's random-forest scoring path:
 *   alg::RandomForest::operator()      /root/reference/code/alg/rf.hxx:35-36
 *   ml::rf::Model::predict             /root/reference/code/ml/rf/rf.hxx:362-372
 *   rf_old::readModelFromBinaryFile    /root/reference/code/ml/rf/ml_rf_model.cxx:459-563
 *   readArray (sparse/dense arrays)    /root/reference/code/ml/rf/ml_rf_model.cxx:48-70
 *
 * PARITY UNPINNED for the traversal: the vote arithmetic lives in the
 * third-party classForest/predictClassTree (ajaiantilal/randomforest-matlab,
 * RF_Class_C/src/classRF.cpp + classTree.cpp; the reference pins no version,
 * /root/reference/README.md:37) whose source is NOT in /root/reference.  This
 * file restates its published algorithm (identical to R randomForest's
 * predictClassTree for numeric splits): start at node 0; while
 * nodestatus[k] != -1 (terminal): m = bestvar[k]-1; go to treemap[2k]-1 if
 * x[m] <= xbestsplit[k] else treemap[2k+1]-1; vote for nodeclass[k].
 * The model-file layout IS pinned: tests read files written here back through
 * the reference's own reader (oracle/_ref/libglia_refrf.so).
 */
#include "glia_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

struct rf_oracle {
  int ntree, nrnodes, nclass, mdim, mtry;
  int* treemap;       /* [ntree][nrnodes][2], 1-based child ids */
  int* nodestatus;    /* [ntree][nrnodes], -1 = terminal */
  double* xbestsplit; /* [ntree][nrnodes] */
  int* bestvar;       /* [ntree][nrnodes], 1-based feature */
  int* nodeclass;     /* [ntree][nrnodes], 1-based class */
  int* ndbigtree;     /* [ntree] */
  int* orig_labels;   /* [nclass] */
};

/* Mirror of rf_old::Model's in-memory layout on x86-64 / libstdc++
 * (ml/rf/ml_rf.h:97-157; sizeof == 520): the file starts with a raw dump of it. */
typedef struct { void* p; int n[2]; } rf_arr;
typedef struct {
  char vec[4][24];          /* four std::vector headers */
  rf_arr ncat, categorical_feature;
  int nrnodes, ntree;
  rf_arr xbestsplit, classwt, cutoff, treemap, nodestatus, nodeclass, bestvar, ndbigtree;
  int mtry, pad0;
  rf_arr orig_labels, new_labels;
  int nclass, pad1;
  rf_arr outcl, outclts, counttr, proximity, proximity_tst, localImp, importance, importanceSD,
      errtr, errts, inbag, votes, oob_times;
} rf_header;

static int read_arr(FILE* f, void** out, size_t esz, int size) {
  *out = NULL;
  if (size <= 0) return 0;
  char* d = (char*)calloc((size_t)size, esz);
  if (size > 128) {
    unsigned char sparse;
    if (fread(&sparse, 1, 1, f) != 1) { free(d); return -1; }
    if (sparse) {
      int num;
      if (fread(&num, 4, 1, f) != 1) { free(d); return -1; }
      for (int i = 0; i < num; ++i) {
        int idx;
        if (fread(&idx, 4, 1, f) != 1 || idx < 0 || idx >= size ||
            fread(d + (size_t)idx * esz, esz, 1, f) != 1) { free(d); return -1; }
      }
      *out = d;
      return 0;
    }
  }
  if (fread(d, esz, (size_t)size, f) != (size_t)size) { free(d); return -1; }
  *out = d;
  return 0;
}

static void write_arr(FILE* f, const void* data, size_t esz, int size, int is_double) {
  if (size <= 0) return;
  if (size > 128) {
    int nnz = 0;
    for (int i = 0; i < size; ++i) {
      double v = is_double ? ((const double*)data)[i] : (double)((const int*)data)[i];
      if (fabs(v) > 1e-8) ++nnz;
    }
    unsigned char sparse = (nnz < size / 2);
    fwrite(&sparse, 1, 1, f);
    if (sparse) {
      fwrite(&nnz, 4, 1, f);
      for (int i = 0; i < size; ++i) {
        double v = is_double ? ((const double*)data)[i] : (double)((const int*)data)[i];
        if (fabs(v) > 1e-8) {
          fwrite(&i, 4, 1, f);
          fwrite((const char*)data + (size_t)i * esz, esz, 1, f);
        }
      }
      return;
    }
  }
  fwrite(data, esz, (size_t)size, f);
}

/* x is rn x cn row-major; returns cn x rn row-major (ml_rf.h:351-362). */
static void* transpose(const void* x, size_t esz, int rn, int cn) {
  char* r = (char*)malloc((size_t)rn * cn * esz);
  for (int i = 0; i < cn; ++i)
    for (int j = 0; j < rn; ++j)
      memcpy(r + ((size_t)i * rn + j) * esz, (const char*)x + ((size_t)j * cn + i) * esz, esz);
  return r;
}

#define NEL(a) ((a).n[0] * (a).n[1])

rf_oracle* rf_oracle_load(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) return NULL;
  rf_header h;
  if (sizeof(h) != 520 || fread(&h, sizeof(h), 1, f) != 1) { fclose(f); return NULL; }
  for (int i = 0; i < 4; ++i) { /* vectors must be empty: begin == end */
    void* b; void* e;
    memcpy(&b, h.vec[i], 8); memcpy(&e, h.vec[i] + 8, 8);
    if (b != e) { fclose(f); return NULL; } /* categorical features unsupported */
  }
  void *ncat = 0, *catf = 0, *xb = 0, *cw = 0, *co = 0, *tm = 0, *ns = 0, *nc = 0, *bv = 0,
       *nb = 0, *ol = 0;
  int bad = 0, tmp;
  bad |= read_arr(f, &ncat, 4, NEL(h.ncat));
  bad |= read_arr(f, &catf, 4, NEL(h.categorical_feature));
  bad |= fread(&tmp, 4, 1, f) != 1; h.nrnodes = tmp;
  bad |= fread(&tmp, 4, 1, f) != 1; h.ntree = tmp;
  bad |= read_arr(f, &xb, 8, NEL(h.xbestsplit));
  bad |= read_arr(f, &cw, 8, NEL(h.classwt));
  bad |= read_arr(f, &co, 8, NEL(h.cutoff));
  bad |= read_arr(f, &tm, 4, NEL(h.treemap));
  bad |= read_arr(f, &ns, 4, NEL(h.nodestatus));
  bad |= read_arr(f, &nc, 4, NEL(h.nodeclass));
  bad |= read_arr(f, &bv, 4, NEL(h.bestvar));
  bad |= read_arr(f, &nb, 4, NEL(h.ndbigtree));
  bad |= fread(&tmp, 4, 1, f) != 1; h.mtry = tmp;
  bad |= read_arr(f, &ol, 4, NEL(h.orig_labels));
  void* nl = 0;
  bad |= read_arr(f, &nl, 4, NEL(h.new_labels));
  bad |= fread(&tmp, 4, 1, f) != 1; h.nclass = tmp;
  fclose(f); /* the remaining arrays (oob statistics) are not needed for scoring */
  free(ncat); free(cw); free(co); free(nl);
  if (catf) { /* any categorical flag set -> unsupported */
    for (int i = 0; i < NEL(h.categorical_feature); ++i) if (((int*)catf)[i]) bad = 1;
    free(catf);
  }
  if (bad || !xb || !tm || !ns || !nc || !bv || !nb || !ol ||
      NEL(h.xbestsplit) != h.nrnodes * h.ntree || NEL(h.treemap) != 2 * h.nrnodes * h.ntree) {
    free(xb); free(tm); free(ns); free(nc); free(bv); free(nb); free(ol);
    return NULL;
  }
  rf_oracle* m = (rf_oracle*)calloc(1, sizeof(*m));
  m->ntree = h.ntree; m->nrnodes = h.nrnodes; m->nclass = h.nclass; m->mtry = h.mtry;
  m->xbestsplit = (double*)transpose(xb, 8, h.xbestsplit.n[0], h.xbestsplit.n[1]);
  m->treemap = (int*)transpose(tm, 4, h.treemap.n[0], h.treemap.n[1]);
  m->nodestatus = (int*)transpose(ns, 4, h.nodestatus.n[0], h.nodestatus.n[1]);
  m->nodeclass = (int*)transpose(nc, 4, h.nodeclass.n[0], h.nodeclass.n[1]);
  m->bestvar = (int*)transpose(bv, 4, h.bestvar.n[0], h.bestvar.n[1]);
  m->ndbigtree = (int*)transpose(nb, 4, h.ndbigtree.n[0], h.ndbigtree.n[1]);
  m->orig_labels = (int*)ol;
  free(xb); free(tm); free(ns); free(nc); free(bv); free(nb);
  int mx = 0;
  for (size_t i = 0; i < (size_t)m->ntree * m->nrnodes; ++i) if (m->bestvar[i] > mx) mx = m->bestvar[i];
  m->mdim = mx;
  return m;
}

static void* dup(const void* p, size_t n) { void* r = malloc(n); memcpy(r, p, n); return r; }

rf_oracle* rf_oracle_from_arrays(int ntree, int nrnodes, int nclass, int mdim, const int* treemap,
                                 const int* nodestatus, const double* xbestsplit,
                                 const int* bestvar, const int* nodeclass, const int* ndbigtree,
                                 const int* orig_labels) {
  rf_oracle* m = (rf_oracle*)calloc(1, sizeof(*m));
  size_t n = (size_t)ntree * nrnodes;
  m->ntree = ntree; m->nrnodes = nrnodes; m->nclass = nclass; m->mdim = mdim; m->mtry = 1;
  m->treemap = (int*)dup(treemap, n * 2 * 4);
  m->nodestatus = (int*)dup(nodestatus, n * 4);
  m->xbestsplit = (double*)dup(xbestsplit, n * 8);
  m->bestvar = (int*)dup(bestvar, n * 4);
  m->nodeclass = (int*)dup(nodeclass, n * 4);
  m->ndbigtree = (int*)dup(ndbigtree, (size_t)ntree * 4);
  m->orig_labels = (int*)dup(orig_labels, (size_t)nclass * 4);
  return m;
}

void rf_oracle_free(rf_oracle* m) {
  if (!m) return;
  free(m->treemap); free(m->nodestatus); free(m->xbestsplit); free(m->bestvar);
  free(m->nodeclass); free(m->ndbigtree); free(m->orig_labels); free(m);
}

int rf_oracle_info(const rf_oracle* m, int* ntree, int* nrnodes, int* nclass, int* mdim) {
  if (!m) return -1;
  if (ntree) *ntree = m->ntree;
  if (nrnodes) *nrnodes = m->nrnodes;
  if (nclass) *nclass = m->nclass;
  if (mdim) *mdim = m->mdim;
  return 0;
}

void rf_oracle_arrays(const rf_oracle* m, int* treemap, int* nodestatus, double* xbestsplit,
                      int* bestvar, int* nodeclass, int* ndbigtree, int* orig_labels) {
  size_t n = (size_t)m->ntree * m->nrnodes;
  if (treemap) memcpy(treemap, m->treemap, n * 8);
  if (nodestatus) memcpy(nodestatus, m->nodestatus, n * 4);
  if (xbestsplit) memcpy(xbestsplit, m->xbestsplit, n * 8);
  if (bestvar) memcpy(bestvar, m->bestvar, n * 4);
  if (nodeclass) memcpy(nodeclass, m->nodeclass, n * 4);
  if (ndbigtree) memcpy(ndbigtree, m->ndbigtree, (size_t)m->ntree * 4);
  if (orig_labels) memcpy(orig_labels, m->orig_labels, (size_t)m->nclass * 4);
}

static void votes1(const rf_oracle* m, const double* x, int* votes) {
  for (int c = 0; c < m->nclass; ++c) votes[c] = 0;
  for (int t = 0; t < m->ntree; ++t) {
    size_t o = (size_t)t * m->nrnodes;
    const int* tm = m->treemap + 2 * o;
    int k = 0;
    while (m->nodestatus[o + k] != -1) {
      int v = m->bestvar[o + k] - 1;
      k = (x[v] <= m->xbestsplit[o + k]) ? tm[2 * k] - 1 : tm[2 * k + 1] - 1;
    }
    votes[m->nodeclass[o + k] - 1] += 1;
  }
}

static int label_index(const rf_oracle* m, int label) {
  for (int i = 0; i < m->nclass; ++i) if (m->orig_labels[i] == label) return i;
  return -1;
}

double rf_oracle_predict1(const rf_oracle* m, const double* x, int D, int label) {
  (void)D;
  int votes[64];
  int li = label_index(m, label);
  if (li < 0 || m->nclass > 64) return -1.0;
  votes1(m, x, votes);
  return votes[li] / (double)m->ntree;
}

/* go_pred_fn / ref_pred_fn adapter: ctx is the rf_oracle*, label -1 = BC_LABEL_MERGE. */
double rf_oracle_pred_cb(const double* row, int dim, void* ctx) {
  return rf_oracle_predict1((const rf_oracle*)ctx, row, dim, -1);
}

int rf_oracle_predict(const rf_oracle* m, const double* X, size_t N, int D, int label, int* votes,
                      double* prob, int threads) {
  int li = label_index(m, label);
  if (li < 0 || m->nclass > 64 || D < m->mdim) return -1;
  long n = (long)N;
#pragma omp parallel for schedule(static) if (threads > 1) num_threads(threads > 1 ? threads : 1)
  for (long i = 0; i < n; ++i) {
    int v[64];
    votes1(m, X + (size_t)i * D, v);
    if (votes) memcpy(votes + (size_t)i * m->nclass, v, (size_t)m->nclass * 4);
    if (prob) prob[i] = v[li] / (double)m->ntree;
  }
  return 0;
}

int rf_oracle_save(const rf_oracle* m, const char* path) {
  FILE* f = fopen(path, "wb");
  if (!f) return -1;
  rf_header h;
  memset(&h, 0, sizeof(h));
  int nt = m->ntree, nr = m->nrnodes;
  h.nrnodes = nr; h.ntree = nt; h.mtry = m->mtry; h.nclass = m->nclass;
  h.xbestsplit.n[0] = nr; h.xbestsplit.n[1] = nt;
  h.treemap.n[0] = nr; h.treemap.n[1] = 2 * nt;
  h.nodestatus.n[0] = nr; h.nodestatus.n[1] = nt;
  h.nodeclass.n[0] = nr; h.nodeclass.n[1] = nt;
  h.bestvar.n[0] = nr; h.bestvar.n[1] = nt;
  h.ndbigtree.n[0] = 1; h.ndbigtree.n[1] = nt;
  h.orig_labels.n[0] = m->nclass; h.orig_labels.n[1] = 1;
  h.new_labels.n[0] = m->nclass; h.new_labels.n[1] = 1;
  fwrite(&h, sizeof(h), 1, f);
  /* file holds each matrix transposed: native [c][r] (c over n[1]) -> file [r][c] */
  double* xb = (double*)transpose(m->xbestsplit, 8, nt, nr);
  int* tm = (int*)transpose(m->treemap, 4, 2 * nt, nr);
  int* ns = (int*)transpose(m->nodestatus, 4, nt, nr);
  int* nc = (int*)transpose(m->nodeclass, 4, nt, nr);
  int* bv = (int*)transpose(m->bestvar, 4, nt, nr);
  fwrite(&nr, 4, 1, f); fwrite(&nt, 4, 1, f);
  write_arr(f, xb, 8, nr * nt, 1);
  write_arr(f, tm, 4, 2 * nr * nt, 0);
  write_arr(f, ns, 4, nr * nt, 0);
  write_arr(f, nc, 4, nr * nt, 0);
  write_arr(f, bv, 4, nr * nt, 0);
  write_arr(f, m->ndbigtree, 4, nt, 0);
  fwrite(&m->mtry, 4, 1, f);
  write_arr(f, m->orig_labels, 4, m->nclass, 0);
  int* nl = (int*)malloc(4 * (size_t)m->nclass);
  for (int i = 0; i < m->nclass; ++i) nl[i] = i + 1;
  write_arr(f, nl, 4, m->nclass, 0);
  fwrite(&m->nclass, 4, 1, f);
  free(xb); free(tm); free(ns); free(nc); free(bv); free(nl);
  fclose(f);
  return 0;
}
