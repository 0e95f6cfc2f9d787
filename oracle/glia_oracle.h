/* TEST INFRASTRUCTURE — not part of the product.
 *
 * CPU restatement (plain C) of GLIA's hmt feature + merge-order hot path, used
 * only as the parity checker for the CUDA path (tests/, __graft_entry__.smoke(),
 * bench.py's cpu_baseline leg).  The product library never links or calls it.
 *
 * Parity status: the RAG / statistics / feature / merge-queue part is PINNED
 * against the unmodified reference headers (oracle/_ref, built by
 * oracle/Makefile from /root/reference/code) by tests/test_oracle_vs_ref.py and
 * by the golden fixtures under tests/golden/.  The random-forest traversal
 * (rf_oracle.c) restates third-party code that is absent from the reference
 * tree (ajaiantilal/randomforest-matlab, no pinned version): that part is
 * "parity unpinned" — see rf_oracle.c.
 */
#ifndef GLIA_ORACLE_H
#define GLIA_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GO_MAX_IMG 4
#define GO_MAX_THR 8

typedef struct go_cfg {
  int ndim;                 /* 2 or 3 (reference: DIMENSION, glia_base.hxx:49-53) */
  int dims[3];              /* x (fastest), y, z */
  int n_rimg, n_rlimg, n_bimg, n_thr;
  const float* rimg[GO_MAX_IMG];  int rbins[GO_MAX_IMG];  double rlo[GO_MAX_IMG],  rhi[GO_MAX_IMG];
  const float* rlimg[GO_MAX_IMG]; int rlbins[GO_MAX_IMG]; double rllo[GO_MAX_IMG], rlhi[GO_MAX_IMG];
  const float* bimg[GO_MAX_IMG];  int bbins[GO_MAX_IMG];  double blo[GO_MAX_IMG],  bhi[GO_MAX_IMG];
  double thr[GO_MAX_THR];
  int normalize;            /* --ns   (main_merge_order_bc.cxx:36-39) */
  int use_log;              /* --logs (main_merge_order_bc.cxx:86-91) */
  int simple;               /* --simpf (bc_feat.hxx:247-279) */
  int use_hist;             /* GLIA_USE_HISTOGRAM_AS_FEATS build (CMakeLists.txt:58-60; type/feat.hxx:608-626) */
  int use_median;           /* GLIA_USE_MEDIAN_AS_FEATS build (CMakeLists.txt:62-64; type/feat.hxx:677-810) */
  int contour_only;         /* TRegionMap(image, mask, true): genContourMap over the image classifies masked-out
                               voxels too (util/struct.hxx:129-146); the mask then hides neighbours only */
} go_cfg;

typedef struct go_state go_state;
typedef struct rf_oracle rf_oracle;
typedef double (*go_pred_fn)(const double* row, int dim, void* ctx);

/* RAG + base statistics from labels (+mask) + pb + images. */
go_state* go_build(const go_cfg* cfg, const uint32_t* labels, const uint32_t* mask,
                   const float* pb);
void go_free(go_state* s);
int go_num_regions(const go_state* s);
int go_num_edges(const go_state* s);       /* directed */
int go_num_candidates(const go_state* s);  /* mutual pairs */
int go_row_dim(const go_state* s, int with_saliency);
/* per-region dumps, ascending label */
void go_get_regions(const go_state* s, uint32_t* label, uint32_t* count, uint32_t* border,
                    int32_t* bbmin /*[R][3]*/, int32_t* bbmax /*[R][3]*/);
/* region image i statistics: hist [R][bins] (u32), sum,sum2 (double), min,max (float) */
void go_get_region_stats(const go_state* s, int kind /*0 rimg, 1 rlimg*/, int i, uint32_t* hist,
                         double* sum, double* sum2, float* mn, float* mx);
/* directed edges ascending (a,b): labels, count, mutual flag, threshold counts [E][T] */
void go_get_edges(const go_state* s, uint32_t* a, uint32_t* b, uint32_t* count, uint8_t* mutual,
                  uint32_t* thr_count);
void go_get_edge_stats(const go_state* s, int i, uint32_t* hist, double* sum, double* sum2,
                       float* mn, float* mx);
/* initial candidates in ascending (min,max) key order; rows may be NULL */
/* swap (optional, [C]): featurise candidate k as (b,a) instead of (a,b) — only matters for
 * equal-area pairs, where the reference's orientation follows unordered_map iteration order. */
void go_init_candidates(go_state* s, uint32_t* keys /*[C][2]*/, double* rows /*[C][dim]*/,
                        const uint8_t* swap);
/* full greedy merge order (struct_merge.hxx:10-33).  Predictor: rf (if non-NULL) else fn.
 * order [M][3], sal [M], rows (optional) [M][dim] = the -b output.  Returns M.
 * If cand_keys/cand_rows non-NULL they receive every featurised candidate in call order
 * (capacity max_cand). n_cand_out = total number featurised. */
int go_greedy(go_state* s, const rf_oracle* rf, go_pred_fn fn, void* ctx, uint32_t* order,
              double* sal, double* rows, uint32_t* cand_keys, double* cand_rows,
              double* cand_scores, size_t max_cand, size_t* n_cand_out,
              const uint8_t* init_swap);
/* bc_feat stage (main_bc_feat.cxx:27-112) for a given order; sal_in may be NULL. */
int go_bc_feat(go_state* s, const uint32_t* order, int n, const double* sal_in, double init_sal,
               double sal_bias, double* rows);

/* SURVEY §8f rank 1.  merge order from the boundary probability alone (hmt/main_merge_order_pb.cxx:14-40):
 * type 1 = approximate median (needs cfg.use_median), 2 = mean; needs bimg[0] == pb.  Returns M. */
int go_merge_order_pb(go_state* s, int type, uint32_t* order, double* sal);
/* pre_merge (gadget/main_pre_merge.cxx:7-87) without the final ITK relabel: order / sal of the conditional greedy run
 * and the label image with merged regions renamed; needs rimg[0] == bimg[0] == pb.  Returns M. */
int go_pre_merge(go_state* s, const int* size_thr, int n_thr, double rpb_thr, uint32_t* order, double* sal,
                 const uint32_t* labels_in, const uint32_t* mask, uint32_t* labels_out);

/* section-classifier features (hmt/main_sc_feat.cxx:27-117) of region pairs across two 2D sections;
 * pairs [n][2] = (label in section 0, label in section 1), rows [n][go_sc_row_dim].  norm_* <= 0: image volume / diagonal. */
int go_sc_row_dim(const go_state* s);
int go_sc_feat(go_state* s0, go_state* s1, const uint32_t* L0, const uint32_t* M0, const uint32_t* L1, const uint32_t* M1,
               const uint32_t* pairs, int n, double norm_area, double norm_len, int use_log, double* rows);

/* ---- random forest (rf_oracle.c) ---- */
rf_oracle* rf_oracle_load(const char* path);  /* rf_old binary model file */
rf_oracle* rf_oracle_from_arrays(int ntree, int nrnodes, int nclass, int mdim,
                                 const int* treemap, const int* nodestatus,
                                 const double* xbestsplit, const int* bestvar,
                                 const int* nodeclass, const int* ndbigtree,
                                 const int* orig_labels);
void rf_oracle_free(rf_oracle* m);
int rf_oracle_info(const rf_oracle* m, int* ntree, int* nrnodes, int* nclass, int* mdim);
void rf_oracle_arrays(const rf_oracle* m, int* treemap, int* nodestatus, double* xbestsplit,
                      int* bestvar, int* nodeclass, int* ndbigtree, int* orig_labels);
/* votes [N][nclass] (int), prob [N] = votes[idx(label)]/ntree.  threads<=0 -> serial. */
int rf_oracle_predict(const rf_oracle* m, const double* X, size_t N, int D, int label,
                      int* votes, double* prob, int threads);
double rf_oracle_pred_cb(const double* row, int dim, void* ctx);
double rf_oracle_predict1(const rf_oracle* m, const double* x, int D, int label);
/* writes the rf_old binary layout (ml_rf_model.cxx:565-…) for fixtures */
int rf_oracle_save(const rf_oracle* m, const char* path);

#ifdef __cplusplus
}
#endif
#endif
