// TEST INFRASTRUCTURE — not part of the product.
// C-ABI wrapper around the UNMODIFIED reference model-file reader/writer
// (ml/rf/ml_rf_model.cxx, compiled in place from /root/reference/code by
// oracle/Makefile).  Used only to pin the rf_old binary layout that
// oracle/rf_oracle.c and the product's glia_b200_rf_load() parse.
#include <cstring>
#include "ml/rf/ml_rf.h"

extern "C" {

// Reads a model through rf_old::readModelFromBinaryFile (ml_rf_model.cxx:459-563)
// and copies the native-layout arrays out.  Pass NULL arrays to query sizes.
int glia_refrf_read(const char* path, int* ntree, int* nrnodes, int* nclass,
                    int* treemap, int* nodestatus, double* xbestsplit, int* bestvar,
                    int* nodeclass, int* ndbigtree, int* orig_labels) {
  rf_old::Model m;
  rf_old::readModelFromBinaryFile(m, path);
  *ntree = m.ntree; *nrnodes = m.nrnodes; *nclass = m.nclass;
  size_t n = (size_t)m.ntree * m.nrnodes;
  if (treemap) std::memcpy(treemap, m.treemap, n * 2 * sizeof(int));
  if (nodestatus) std::memcpy(nodestatus, m.nodestatus, n * sizeof(int));
  if (xbestsplit) std::memcpy(xbestsplit, m.xbestsplit, n * sizeof(double));
  if (bestvar) std::memcpy(bestvar, m.bestvar, n * sizeof(int));
  if (nodeclass) std::memcpy(nodeclass, m.nodeclass, n * sizeof(int));
  if (ndbigtree) std::memcpy(ndbigtree, m.ndbigtree, m.ntree * sizeof(int));
  if (orig_labels) std::memcpy(orig_labels, m.orig_labels, m.nclass * sizeof(int));
  return 0;
}

// Round trip through the reference writer (ml_rf_model.cxx writeModelToBinaryFile).
int glia_refrf_rewrite(const char* in_path, const char* out_path) {
  rf_old::Model m;
  rf_old::readModelFromBinaryFile(m, in_path);
  rf_old::writeModelToBinaryFile(out_path, m);
  return 0;
}

// Writes a model the way training does: ml_rf_train.cxx:696-719 stores each native classRF
// array TRANSPOSED in the Model (dims swapped) and writeModelToBinaryFile dumps it as is.
// (A Model that came from readModelFromBinaryFile holds native arrays, so re-writing it is not
// a supported round trip in the reference.)  Native: [ntree][nrnodes] (treemap [ntree][2*nrnodes]).
int glia_refrf_write_native(const char* path, int ntree, int nrnodes, int nclass, int mtry,
                            int* treemap, int* nodestatus, double* xbestsplit, int* bestvar,
                            int* nodeclass, int* ndbigtree, int* orig_labels) {
  rf_old::Model m;
  m.nrnodes = nrnodes; m.ntree = ntree; m.mtry = mtry; m.nclass = nclass;
  m.xbestsplit = rf_old::transpose(xbestsplit, ntree, nrnodes);
  m.n_xbestsplit[0] = nrnodes; m.n_xbestsplit[1] = ntree;
  m.treemap = rf_old::transpose(treemap, 2 * ntree, nrnodes);
  m.n_treemap[0] = nrnodes; m.n_treemap[1] = 2 * ntree;
  m.nodestatus = rf_old::transpose(nodestatus, ntree, nrnodes);
  m.n_nodestatus[0] = nrnodes; m.n_nodestatus[1] = ntree;
  m.nodeclass = rf_old::transpose(nodeclass, ntree, nrnodes);
  m.n_nodeclass[0] = nrnodes; m.n_nodeclass[1] = ntree;
  m.bestvar = rf_old::transpose(bestvar, ntree, nrnodes);
  m.n_bestvar[0] = nrnodes; m.n_bestvar[1] = ntree;
  m.ndbigtree = rf_old::transpose(ndbigtree, ntree, 1);
  m.n_ndbigtree[0] = 1; m.n_ndbigtree[1] = ntree;
  m.orig_labels = new int[nclass];  // ~Model deletes its arrays
  std::memcpy(m.orig_labels, orig_labels, nclass * sizeof(int));
  m.n_orig_labels[0] = nclass; m.n_orig_labels[1] = 1;
  rf_old::writeModelToBinaryFile(path, m);
  return 0;
}

int glia_refrf_sizeof_model() { return (int)sizeof(rf_old::Model); }

}  // extern "C"
