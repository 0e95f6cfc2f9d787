/* glia_b200 — C ABI of the B200-native hmt feature + merge-order path.
 *
 * The reference (tingliu/glia) has no FFI layer; the seams this library sits behind are
 * (reference paths relative to /root/reference/code):
 *   - TRegionMap(seg, mask, false)                     type/region_map.hxx:52-95
 *     genPointMap / genContourMap / getContourTraits   util/struct.hxx:61-144, type/neighbor.hxx:72-126
 *       -> glia_b200_build()  (RAG + per-region / per-directed-boundary statistics)
 *   - TBoundaryTable::init + fBcFeat + fBcPred         type/boundary_table.hxx:91-114,
 *                                                      hmt/main_merge_order_bc.cxx:54-139
 *       -> glia_b200_init_candidates(), glia_b200_score_candidates()
 *   - genMergeOrderGreedyUsingBoundaryClassifier       util/struct_merge_bc.hxx:8-58
 *       -> glia_b200_merge_order_bc()
 *   - bc_feat operation()                              hmt/main_bc_feat.cxx:27-112
 *       -> glia_b200_bc_feat()
 *   - alg::RandomForest::operator(), Model::predict    alg/rf.hxx:35-36, ml/rf/rf.hxx:362-405
 *     extern classForest(...)                          ml/rf/rf.hxx:8-14
 *     rf_old::readModelFromBinaryFile                  ml/rf/ml_rf_model.cxx:459-563
 *       -> glia_b200_rf_load(), glia_b200_rf_from_arrays(), glia_b200_rf_score()
 *
 * Conventions: every function returns an int status (GLIA_OK == 0); nothing calls exit().
 * The caller owns every buffer passed in or out.  Buffers are HOST pointers unless the
 * handle was created with device_pointers != 0, in which case the bulk inputs (labels, mask,
 * images, RF feature matrix) and bulk outputs documented as "host or device" are DEVICE
 * pointers on the handle's GPU.  One handle per GPU/stream; a handle is not thread-safe.
 * There is no CPU fallback: every entry point fails with GLIA_E_CUDA when no device is usable.
 */
#ifndef GLIA_B200_H
#define GLIA_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GLIA_API __attribute__((visibility("default")))
#else
#define GLIA_API
#endif

#define GLIA_MAX_IMG 4
#define GLIA_MAX_THR 8
#define GLIA_MAX_BINS 64

enum {
  GLIA_OK = 0,
  GLIA_E_ARG = 1,      /* bad argument / unsupported configuration */
  GLIA_E_CUDA = 2,     /* CUDA runtime error (see glia_b200_last_error) */
  GLIA_E_CAPACITY = 3, /* edge table too small: recreate with a larger edge_capacity */
  GLIA_E_STATE = 4,    /* call order violated (e.g. candidates before build) */
  GLIA_E_IO = 5,       /* model file unreadable / malformed */
  GLIA_E_ORDER = 6     /* merge order references an unknown region */
};

/* One image + histogram spec: ImageHistPair (type/feat.hxx) = --rbi/--rbb/--rbl/--rbu etc. */
typedef struct glia_image_spec {
  const float* data; /* same extent as the label volume */
  int bins;          /* 1..GLIA_MAX_BINS */
  double lo, hi;     /* histogram range (util/image_stats.hxx:12-36 semantics) */
} glia_image_spec;

typedef struct glia_params {
  int ndim;    /* 2 or 3: the reference's compile-time DIMENSION (glia_base.hxx:49-53) */
  int dims[3]; /* x (fastest), y, z; dims[2] ignored when ndim == 2 */
  int n_rimg, n_rlimg, n_bimg, n_thr;
  glia_image_spec rimg[GLIA_MAX_IMG];  /* region images   (rImages,  hmt_util.hxx:17-56) */
  glia_image_spec rlimg[GLIA_MAX_IMG]; /* label images    (rlImages) */
  glia_image_spec bimg[GLIA_MAX_IMG];  /* boundary images (bImages) */
  double thr[GLIA_MAX_THR];            /* --bt thresholds on pb (type/feat.hxx:485-502) */
  int normalize;                       /* --ns   (main_merge_order_bc.cxx:36-39) */
  int use_log;                         /* --logs (main_merge_order_bc.cxx:86-91) */
  int simple;                          /* --simpf (hmt/bc_feat.hxx:247-279) */
  int device_pointers;                 /* 0: host buffers, 1: device buffers */
  int device;                          /* CUDA device ordinal */
  uint32_t label_bound;  /* exclusive upper bound of label values; 0 = scan the volume */
  uint64_t edge_capacity; /* slots of the directed-boundary table; 0 = automatic */
  /* z-slab sharding (multi-GPU).  The buffers given to glia_b200_build hold
   * planes [z_begin - halo_lo, z_end + halo_hi) of a volume whose full extent is dims[];
   * only voxels of planes [z_begin, z_end) are owned (classified and accumulated), the halo
   * planes supply neighbour labels/mask.  z_end == 0 means the whole volume (no slab). */
  int z_begin, z_end, halo_lo, halo_hi;
  void* stream; /* cudaStream_t to run on (NULL = a private stream) */
  /* The reference's two compile-time feature layouts (CMakeLists.txt:53-64) as run-time fields:
   * use_hist_feat   = GLIA_USE_HISTOGRAM_AS_FEATS: every ImageLabelFeats block is [histogram bins] + entropy
   *                   (type/feat.hxx:608-626);
   * use_median_feat = GLIA_USE_MEDIAN_AS_FEATS: every ImageRealFeats block is median, mean, stddev, min, max, the
   *                   difference blocks gain |median difference| (type/feat.hxx:677-700, :772-810) and --simpf
   *                   rows the pair boundary's median (hmt/bc_feat.hxx:252-268).  median = the n/2-th order
   *                   statistic (util/stats.hxx:83-91).  Needs the whole volume in one handle (no z-slab). */
  int use_hist_feat, use_median_feat;
  /* TRegionMap(image, mask, onlyContour = true) (type/region_map.hxx:54-58), as hmt/main_merge_order_pb.cxx:23 builds
   * it: the image-driven genContourMap (util/struct.hxx:129-146) classifies EVERY voxel, masked-out ones included -
   * the mask then only hides neighbours.  Region records count those voxels too; only boundary records are
   * meaningful in this mode (which is all glia_b200_merge_order_pb reads). */
  int contour_only;
} glia_params;

typedef struct glia_handle glia_handle;
typedef struct glia_rf glia_rf;

GLIA_API int glia_b200_create(const glia_params* p, glia_handle** out);
GLIA_API void glia_b200_destroy(glia_handle* h);
GLIA_API const char* glia_b200_last_error(const glia_handle* h);
/* Re-point a configured image (kind 0: rimg, 1: rlimg, 2: bimg; index i) at another buffer of the same
 * extent before the next glia_b200_build: one handle serves a stream of slices / volumes, as the
 * reference's per-image process loop does (gadget/python/script_util.py:6-23). */
GLIA_API int glia_b200_set_image(glia_handle* h, int kind, int i, const float* data);

/* SURVEY section 8f rank 1.  Merge order from the boundary probability alone (hmt/main_merge_order_pb.cxx:14-40,
 * util/struct_merge.hxx:36-136): saliency of a pair = -(mean | approximate median) of pb over its pair boundary, a merge
 * combines the data of the items it replaces (weighted mean, or concatenated values) and never re-reads the image.
 * type: 1 = median (the handle must have use_median_feat), 2 = mean.  Boundary image 0 must be the pb map.  The
 * reference builds its region map with onlyContour = true: create the handle with contour_only = 1 to reproduce it
 * when a mask is given.  order [R-1][3] / sal [R-1] are host buffers; *n_merges receives the number of merges. */
GLIA_API int glia_b200_merge_order_pb(glia_handle* h, int type, uint32_t* order, double* sal, int64_t* n_merges);
/* pre_merge (gadget/main_pre_merge.cxx:7-87): the mean-pb greedy chain restricted to pairs whose smaller region is
 * below size_thresholds[0] voxels, or - with a second threshold - where a region below size_thresholds[1] has a mean
 * pb above rpb_threshold; regions are updated as they merge.  Region image 0 and boundary image 0 must be the pb map.
 * labels_out (optional; host or device as the handle's inputs, same extent as the label volume) receives the label
 * volume with every merged region renamed to its final key (transformKeys + transformImage); the reference's
 * optional ITK relabel step is not part of this call. */
GLIA_API int glia_b200_pre_merge(glia_handle* h, const int* size_thresholds, int n_thresholds, double rpb_threshold,
                                 uint32_t* order, double* sal, int64_t* n_merges, uint32_t* labels_out);

/* SURVEY section 8f rank 3.  segment_greedy (hmt/main_segment_greedy.cxx:30-95): merge trees (one per order) with node
 * potentials from the merge probabilities (hmt/tree_build.hxx:41-62; all 1 when merge_probs is NULL), optionally scaled
 * by per-node region probabilities in tree node order, are resolved greedily over all trees (hmt/tree_greedy.hxx:107-152:
 * repeatedly the valid node of largest potential, which invalidates its ancestors, its descendants and - through its
 * leaf labels - the matching leaves and their ancestors in the other trees).  Pick k renames the superpixels below it
 * to k + 1 (hmt/tree_segment.hxx:24-35); labels_out (host or device as the handle's inputs) receives the volume the
 * handle was built from with those names, unmapped superpixels set to 0 when ignore_missing (the reference's default)
 * and left alone otherwise, masked-out voxels untouched.  picks (optional, [2 R][2]) = (tree, node) in pick order with
 * nodes numbered as genTree numbers them (first appearance in the order).  The boundary-confidence image (-b) and the
 * ITK relabel (-r) of the reference program are not part of this call. */
GLIA_API int glia_b200_segment_greedy(glia_handle* h, int n_trees, const uint32_t* const* orders, const int64_t* n_merges,
                                      const double* const* merge_probs, const double* const* region_probs,
                                      int ignore_missing, uint32_t* labels_out, int32_t* picks, int64_t* n_picks);

/* SURVEY section 8f rank 2.  Boundary-classifier training labels from a ground-truth segmentation
 * (hmt/main_bc_label_ri.cxx:31-155; hmt/bc_label.hxx:44-122; util/image_stats.hxx:173-245; util/stats.hxx:186-265):
 * for every merge of `order` decide BC_LABEL_MERGE (-1) or BC_LABEL_SPLIT (+1) by comparing the pixel-pair F1 score
 * (use_pair_f1, the reference's default) or the Rand index of the merged region against that of the two parts, counted
 * against `truth` (same extent as the label volume, host or device as the handle's inputs; truth label 0 is background
 * and excluded; masked-out voxels are in no region).  tweak / max_prec_drop: the reference's -w / -d rules;
 * opt_split: compare against the best split of the two sides found so far (-p); global_opt 0: every merge on its own,
 * 1 / 2: best splits bottom-up, then path consistency by merging / by splitting (-g).  bc_labels [n] (host);
 * merge_score / split_score [n] (host, optional): the F1 or Rand index values behind the labels (global_opt 0). */
GLIA_API int glia_b200_bc_label(glia_handle* h, const uint32_t* truth, const uint32_t* order, int64_t n,
                                int use_pair_f1, int opt_split, int global_opt, int tweak, double max_prec_drop,
                                int32_t* bc_labels, double* merge_score, double* split_score);

/* SURVEY section 8f rank 4.  Section-classifier features (hmt/main_sc_feat.cxx:27-117; hmt/sc_feat.hxx): one row per listed
 * pair (region of section 0, region of section 1) across two 2D sections of equal extent, each built into its own
 * handle (same device, same feature configuration; not with use_median_feat / simple).  Row = RegionPairFeats
 * (shape, centroid distance, overlap / set differences, central-moment, Hu-moment and eccentricity differences, image
 * differences) | RegionFeatsWithLocation of the region with the smaller area | that of the larger one (RegionFeats +
 * 7 central moments, 7 Hu moments, eccentricity: type/feat.hxx:191-268, alg/geometry.hxx:44-108).  pairs [n][2] =
 * labels (host); norm_area / norm_length <= 0 = image volume / diagonal (main_sc_feat.cxx:48-56); use_log = --logs.
 * rows [n][glia_b200_sc_row_dim] (host, or device when h0 was created with device_pointers). */
GLIA_API int glia_b200_sc_row_dim(const glia_handle* h);
GLIA_API int glia_b200_sc_feat(glia_handle* h0, glia_handle* h1, const uint32_t* pairs, int64_t n, double norm_area,
                               double norm_length, int use_log, double* rows);

/* Per-handle knobs (nothing on the data path reads the environment).  key "generic": value != 0 forces the
 * generic voxel kernel (test hook: fast == generic); "tz": planes (2D: rows) per tile of the fast voxel
 * kernel, "slots": its shared-memory table slots (0 = default); "notma": value != 0 makes the generic voxel kernel
 * stage its label tile with plain loads instead of a TMA tensor copy; "used_tma" (query, value ignored): GLIA_OK when
 * the last generic pass staged by TMA, GLIA_E_STATE otherwise. */
GLIA_API int glia_b200_tune(glia_handle* h, const char* key, int value);

/* RAG + base statistics.  labels/mask: uint32 volumes (mask may be NULL); pb: float32. */
GLIA_API int glia_b200_build(glia_handle* h, const uint32_t* labels, const uint32_t* mask, const float* pb);
/* R regions, E directed boundaries, C mutual pairs (= initial merge candidates) */
GLIA_API int glia_b200_counts(const glia_handle* h, int64_t* R, int64_t* E, int64_t* C);
GLIA_API int glia_b200_row_dim(const glia_handle* h, int with_saliency);

/* Dumps (host buffers; any pointer may be NULL).  Regions ascending by label. */
GLIA_API int glia_b200_get_regions(glia_handle* h, uint32_t* label, uint32_t* count, uint32_t* border,
                          int32_t* bbmin /*[R][3]*/, int32_t* bbmax /*[R][3]*/);
/* kind 0: rimg[i], kind 1: rlimg[i].  hist [R][bins]; min/max are 0 for empty sets. */
GLIA_API int glia_b200_get_region_stats(glia_handle* h, int kind, int i, uint32_t* hist, double* sum,
                               double* sum2, float* mn, float* mx);
/* Directed boundaries ascending by (a, b) label. thr_count [E][n_thr]. */
GLIA_API int glia_b200_get_edges(glia_handle* h, uint32_t* a, uint32_t* b, uint32_t* count,
                        uint8_t* mutual, uint32_t* thr_count);
GLIA_API int glia_b200_get_edge_stats(glia_handle* h, int i, uint32_t* hist, double* sum, double* sum2,
                             float* mn, float* mx);

/* Initial candidates in ascending (min,max) label order: keys [C][2], rows [C][dim]
 * (host or device; either may be NULL).  Candidate (a,b), a<b, is featurised as
 * fBcFeat(reg a, reg b): for equal areas x1 = a unless set_tie_orientation says otherwise (below). */
GLIA_API int glia_b200_init_candidates(glia_handle* h, uint32_t* keys, double* rows);
/* Equal-area candidates.  fBcFeat swaps its two regions on a strict ">" of the areas
 * (hmt/main_merge_order_bc.cxx:76-80), so for EQUAL areas x1 is the region TBoundaryTable::init passed first,
 * i.e. the one its loop over the std::unordered_map of regions meets first (type/boundary_table.hxx:92-107) -
 * a property of the C++ library, not of the algorithm.  Default here: x1 = smaller label.
 *   candidate_ties: tie[c] = 1 for initial candidates with equal areas (host buffer, may be NULL), *n_ties = count:
 *     the complete list of rows that can differ from a reference run by an x1/x2 block swap.
 *   set_tie_orientation: second_first[c] != 0 makes the larger label x1 when the areas of candidate c tie
 *     (host buffer of C bytes; NULL restores the default).  Applies to init_candidates, score_candidates and
 *     merge_order_bc until the next build.
 *   reference_orientation: fills second_first[C] with the call order of a reference binary built with this
 *     toolchain's libstdc++, by replaying the reference's three chained unordered_maps (count map in raster
 *     order of first appearance -> point map -> region map, util/struct.hxx:61-92, type/region_map.hxx:79-95)
 *     and its loop, including the scratch-region insertion of util/struct_merge_bc.hxx:19-20.  reserve_hint > 0
 *     applies rmap.reserve(reserve_hint) first (harnesses that make the loop rehash-free do that);
 *     0 = as hmt/main_merge_order_bc.cxx runs.  Needs the labels (and mask) of the last build still in place;
 *     whole volumes only. */
GLIA_API int glia_b200_candidate_ties(glia_handle* h, uint8_t* tie, int64_t* n_ties);
GLIA_API int glia_b200_set_tie_orientation(glia_handle* h, const uint8_t* second_first);
GLIA_API int glia_b200_reference_orientation(glia_handle* h, int64_t reserve_hint, uint8_t* second_first);

/* Featurise + RF-score the initial candidates on the device: sal [C] (host or device). */
GLIA_API int glia_b200_score_candidates(glia_handle* h, const glia_rf* rf, double* sal);

/* Full greedy merge order (hmt/main_merge_order_bc.cxx:143-157).  order [R-1][3], sal [R-1],
 * rows [R-1][dim] or NULL (the -b output).  *n_merges receives the number of merges. */
GLIA_API int glia_b200_merge_order_bc(glia_handle* h, const glia_rf* rf, uint32_t* order, double* sal,
                             double* rows, int64_t* n_merges, int64_t* n_featurized);
/* bc_feat stage for a given order (hmt/main_bc_feat.cxx:27-112).  sal_in NULL = no saliency
 * features; rows [n][row_dim(sal_in != NULL)]. */
GLIA_API int glia_b200_bc_feat(glia_handle* h, const uint32_t* order /*[n][3]*/, int64_t n,
                      const double* sal_in, double init_sal, double sal_bias, double* rows);

/* ---- multi-GPU: z-slab sharding (no counterpart in the reference, which is single-process;
 * preserves its results because every voxel is owned by exactly one slab).  One process and one
 * handle per GPU, created with z_begin/z_end/halo_* and a label_bound common to all ranks.  Call
 * order (glia_b200/parallel.py drives it with NCCL collectives):
 *   build_partial -> all-reduce raw_region_tables (add: sum, max: unsigned max, f64: sum)
 *   -> finalize_regions -> edge_interest (all-reduce max over the byte mask)
 *   -> export_edges -> all-to-all of the per-destination record runs -> import_edges
 *   -> finalize_edges(own_lo, own_hi) -> init_candidates / score_candidates on the owned candidates.
 * Region index r (ascending label order, identical on all ranks) is owned by rank
 * min(world-1, r / ceil(R/world)); a candidate (a,b), a<b, belongs to the owner of a.
 * All returned pointers are DEVICE pointers owned by the handle. */
GLIA_API int glia_b200_build_partial(glia_handle* h, const uint32_t* labels, const uint32_t* mask, const float* pb);
GLIA_API int glia_b200_raw_region_tables(glia_handle* h, void** add, void** mx, void** f64, int64_t* n_add,
                                         int64_t* n_max, int64_t* n_f64);
GLIA_API int glia_b200_finalize_regions(glia_handle* h, int64_t* R);
/* The same in three steps, for callers that overlap the all-reduce of the record tables with the exchange of
 * the boundary records: region_presence returns the u32 presence flags [label_bound] of the LOCAL table
 * (all-reduce max them), index_regions turns the reduced flags into the dense region index (R), and
 * gather_regions compacts the records once the tables of raw_region_tables are reduced.  edge_interest,
 * export_edges, import_edges and finalize_edges need only the index, not the records. */
GLIA_API int glia_b200_region_presence(glia_handle* h, void** flags, int64_t* n);
GLIA_API int glia_b200_index_regions(glia_handle* h, int64_t* R);
GLIA_API int glia_b200_gather_regions(glia_handle* h);
GLIA_API int glia_b200_edge_interest(glia_handle* h, int world, void** mask, int64_t* bytes);
/* counts [world]: records for each destination rank, stored back to back in *buf in rank order.
 * self_rank >= 0: nothing is exported to that rank - its records stay in the handle's table and the
 * received ones are merged with import_edges(clear_first = 0); self_rank < 0: every destination gets a run
 * and the table is rebuilt with import_edges(clear_first = 1). */
GLIA_API int glia_b200_export_edges(glia_handle* h, int world, int self_rank, int64_t* counts, void** buf,
                                    int* rec_bytes);
/* occupied slots (counted by the last export_edges) and capacity of the boundary table: a caller that
 * keeps its own records (self_rank >= 0) checks that occupied + received fits before import_edges(…, 0),
 * and otherwise repeats the exchange with self_rank = -1 / clear_first = 1, which grows the table. */
GLIA_API int glia_b200_table_load(const glia_handle* h, int64_t* occupied, int64_t* capacity);
GLIA_API int glia_b200_import_edges(glia_handle* h, const void* buf, int64_t n, int clear_first);
GLIA_API int glia_b200_finalize_edges(glia_handle* h, int64_t own_lo, int64_t own_hi);

/* The same sequence as ONE call over NCCL, stream-ordered on the handle's stream (shard.cu): the host waits only where
 * it needs a number to size a buffer.  NCCL is bound at run time (dlopen of libnccl.so.2), the library has no link-time
 * dependency on it.  glia_b200_comm_unique_id fills 256 bytes (two ncclUniqueId: boundary exchange / record-table
 * all-reduce) on one rank; the caller hands them to every rank (any transport), which then calls glia_b200_comm_create
 * collectively.  One communicator per handle / GPU; ranks may be processes or threads of one process. */
typedef struct glia_comm glia_comm;
GLIA_API int glia_b200_comm_unique_id(void* id256);
GLIA_API int glia_b200_comm_create(const void* id256, int world, int rank, int device, glia_comm** out);
GLIA_API void glia_b200_comm_destroy(glia_comm* c);
GLIA_API int glia_b200_build_sharded(glia_handle* h, glia_comm* c, const uint32_t* labels, const uint32_t* mask,
                                     const float* pb);
/* The bc_feat stage (glia_b200_bc_feat) on a handle built by glia_b200_build_sharded: collective over the communicator.
 * Every rank passes the same order / saliencies; rank r featurises merges [*first, *first + *count) - contiguous
 * shares of ceil(n / world) - and writes them to rows[0 .. *count) (reference: the parfor over merges of
 * hmt/main_bc_feat.cxx:76-95).  The per-node boundary tables are completed by all-reduces, the merge tree is
 * replicated.  Not available with use_median_feat. */
GLIA_API int glia_b200_bc_feat_sharded(glia_handle* h, glia_comm* c, const uint32_t* order, int64_t n,
                                       const double* sal_in, double init_sal, double sal_bias, double* rows,
                                       int64_t* first, int64_t* count);
GLIA_API double glia_b200_comm_last_exchange_ms(const glia_comm* c);

/* ---- random forest ---- */
GLIA_API int glia_b200_rf_load(const char* path, int device, glia_rf** out);
/* Native layouts after readModelFromBinaryFile's transposes (ml_rf_model.cxx:542-557):
 * treemap [ntree][nrnodes][2] (1-based child ids), others [ntree][nrnodes]. */
GLIA_API int glia_b200_rf_from_arrays(int ntree, int nrnodes, int nclass, int mdim, const int* treemap,
                             const int* nodestatus, const double* xbestsplit, const int* bestvar,
                             const int* nodeclass, const int* ndbigtree, const int* orig_labels,
                             int device, glia_rf** out);
/* alg::EnsembleRandomForest with opt::ThresholdModelDistributor (alg/rf.hxx:63-98,
 * type/function.hxx:71-85; --bcm m0 m1 m2 --bcmd dim0 dim1 thr): a row x is scored by model 0 if
 * x[dim1] < threshold, else by model 1 if x[dim0] < threshold, else by model 2.  n = 1..3 models on one
 * device with identical class labels; the result is an independent forest (the inputs may be destroyed). */
GLIA_API int glia_b200_rf_ensemble(const glia_rf* const* models, int n, int dim0, int dim1, double threshold,
                                   glia_rf** out);
GLIA_API void glia_b200_rf_destroy(glia_rf* m);
GLIA_API int glia_b200_rf_info(const glia_rf* m, int* ntree, int* nrnodes, int* nclass, int* mdim);
/* X [N][D] row-major doubles (x[m + i*mdim], ml/rf/rf.hxx:362-365).  votes [N][nclass] int
 * (may be NULL), prob [N] = votes[idx(label)]/ntree (ml/rf/rf.hxx:366-369).
 * device_pointers: X, votes, prob are device buffers. stream may be NULL. */
GLIA_API int glia_b200_rf_score(const glia_rf* m, const double* X, size_t N, int D, int label, int* votes,
                       double* prob, int device_pointers, void* stream);
/* Allocation-free variant of glia_b200_rf_score (prob only, DEVICE buffers).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The per-row vote table [N][nclass] that the vote-counting scorer needs as scratch
 * lives in a static __device__ array of GLIA_NOALLOC_MAX_VOTES (1 Mi) int counters = 4 MiB, held by every GPU that loads
 * the library until the process exits; the call itself runs no cudaMalloc / cudaFree / allocating library routine.
 * d_rows [N][D] and d_prob [N] are device buffers on the forest's device; the work is queued on `stream` (may be NULL)
 * and d_prob is complete when that stream has drained.  N * nclass > GLIA_NOALLOC_MAX_VOTES is rejected with
 * GLIA_E_CAPACITY and a message in err (optional buffer of err_len bytes).  All calls on a device share the one
 * array: each call clears the counters it uses first, and calls that overlap on the GPU (different streams with no
 * ordering between them) race on it - serialise them.  Same votes, hence the same probabilities, as glia_b200_rf_score. */
#define GLIA_NOALLOC_MAX_VOTES (1u << 20)
GLIA_API int glia_b200_rf_score_noalloc(const glia_rf* m, const double* d_rows, size_t N, int D, int label,
                                        double* d_prob, void* stream, char* err, size_t err_len);
GLIA_API size_t glia_b200_rf_score_noalloc_max_votes(void);

/* Timing of the last glia_b200_build/score call, measured with CUDA events on the handle's
 * stream: which = 0 voxel-pass kernel, 1 graph (compaction+sort+CSR), 2 featurise, 3 RF score,
 * 4 table initialisation (+ label scan when label_bound == 0).
 * Returns milliseconds, or a negative value if not available. */
GLIA_API double glia_b200_last_ms(const glia_handle* h, int which);
/* Number of kernels launched by this handle so far (bench.py's gpu_launches). */
GLIA_API int64_t glia_b200_launch_count(const glia_handle* h);

/* Bench/test utility (not part of the reference's path): seeded jittered-Voronoi superpixels +
 * dyadic boundary map written straight into DEVICE buffers; same integer hashing as
 * glia_b200/synth.py::voronoi_volume.  dims: x (fastest), y, z. */
GLIA_API int glia_b200_synth_voronoi(int ndim, const int* dims, int cell, uint32_t seed,
                                     uint32_t* labels, float* pb, void* stream, uint64_t* n_cells);
/* planes [z0, z0+nz) of the same 3D volume: one rank's z-slab with its halo */
GLIA_API int glia_b200_synth_voronoi_slab(const int* dims, int cell, uint32_t seed, int z0, int nz,
                                          uint32_t* labels, float* pb, void* stream, uint64_t* n_cells);

#ifdef __cplusplus
}
#endif
#endif
