/* orbx_bow.h — C-ABI for SURVEY.md §8f N2: Frame::ComputeBoW (/root/reference/src/Frame.cc:447-454), i.e.
 * ORBVocabulary::transform(features, BowVector&, FeatureVector&, levelsup = 4)
 * (/root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1207, 1231-1272) in the configuration ORB-SLAM2
 * ships: weighting TF_IDF, scoring L1_NORM, FORB descriptors (256-bit, Hamming).
 *
 * Same conventions as orbx.h (status codes, orbx_last_error(), host pointers, no CPU fallback). Part of liborbx.so.
 */
#ifndef ORBX_BOW_H
#define ORBX_BOW_H
#include "orbx.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbx_vocabulary orbx_vocabulary;

/* The vocabulary as TemplatedVocabulary::loadFromTextFile builds it (TemplatedVocabulary.h:1351-1436): entry i describes
 * node id i+1 (the root is node 0) — parent id (< i+1), leaf flag, 32-byte descriptor, weight. Children keep their order
 * of appearance; word ids are assigned to leaves in order of appearance. L = m_L (depth). A node must be a leaf exactly
 * when it has no children, and no node may have more than 32 children (the loader allows k <= 20). */
int orbx_vocabulary_create(int device, int L, int n_nodes, const int* parent, const uint8_t* is_leaf, const uint8_t* desc,
                           const double* weight, orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* v);
int orbx_vocabulary_words(const orbx_vocabulary* v);

/* transform() for n_frames descriptor sets in one call. desc[f]: n[f] x 32 bytes (mDescriptors).
 * BowVector of frame f -> word_ids[f] / word_vals[f] ascending by word id, n_words[f] entries (capacity n[f]);
 * FeatureVector -> CSR ascending by node id: fv_nodes[f][m], fv_start[f][m+1], fv_feats[f] (capacities n[f], n[f]+1,
 * n[f]); n_fv_nodes[f] = m. Values are the reference's doubles bit for bit (additions in feature order, the L1 norm
 * summed in word order). */
int orbx_compute_bow(orbx_vocabulary* v, const uint8_t* const* desc, const int* n, int n_frames, int levelsup, int* const* word_ids,
                     double* const* word_vals, int* n_words, int* const* fv_nodes, int* const* fv_start, int* const* fv_feats,
                     int* n_fv_nodes);

/* The same on descriptors that are still on the device: frames first_frame .. first_frame + n_frames - 1, side `side`
 * (0 = left / mono) of a batch engine after orbx_batch_run / orbx_batch_process* — as Frame::ComputeBoW works on the
 * mDescriptors the extractor just produced (src/Frame.cc:447-454), without carrying them through the host. n[f] = the
 * frame's keypoint count (as downloaded). Outputs and values identical to orbx_compute_bow. */
int orbx_compute_bow_resident(orbx_vocabulary* v, orbx_batch* batch, int first_frame, int side, const int* n, int n_frames,
                              int levelsup, int* const* word_ids, double* const* word_vals, int* n_words, int* const* fv_nodes,
                              int* const* fv_start, int* const* fv_feats, int* n_fv_nodes);

#ifdef __cplusplus
}
#endif
#endif
