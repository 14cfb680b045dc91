/* sc_c_api.h -- C ABI of the ScanContext place-recognition path on the device, exported
 * by the same libndt_b200.so as ndt_c_api.h.  It replaces, for many scans / pairs / queries per call, the members of the
 * reference's SCManager (xchu_mapping/include/scancontext/Scancontext.h:56-121, Scancontext.cpp = SC):
 *
 *   ndtb_sc_add_scans   SCManager::makeAndSaveScancontextAndKeys (SC:249-260): makeScancontext (SC:151-196),
 *                       makeRingkeyFromScancontext / makeSectorkeyFromScancontext (SC:199-228), eig2stdvec (SC:64-68)
 *   ndtb_sc_distance    SCManager::distanceBtnScanContext (SC:118-148): fastAlignUsingVkey (SC:96-115), distDirectSC (SC:71-93)
 *   ndtb_sc_detect      SCManager::detectLoopClosureID (SC:331-423) for given query descriptors and tree sizes: exact k nearest
 *                       ring keys (the reference: nanoflann kd-tree, L2, eps 0), column-shift distance to each, threshold
 *
 * Descriptors, keys and the ring-key matrix stay resident in HBM.  All arithmetic is the reference's (double descriptors, float
 * ring keys for the search); results are bit-identical to the CPU oracle's (tests/test_scancontext.py).  Plain pointers and sizes
 * only; every function returns 0 on success, non-zero on failure with the reason in ndtb_sc_last_error.  No CPU fallback. */
#ifndef NDT_B200_SC_C_API_H_
#define NDT_B200_SC_C_API_H_
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ndtb_sc ndtb_sc;

/* Scancontext.h:83-103 */
typedef struct ndtb_sc_params {
  int num_ring;            /* PC_NUM_RING 20 (<= 64) */
  int num_sector;          /* PC_NUM_SECTOR 60 (<= 128) */
  double max_radius;       /* PC_MAX_RADIUS 80.0 */
  double lidar_height;     /* LIDAR_HEIGHT 2.0 */
  double search_ratio;     /* SEARCH_RATIO 0.1 */
  int num_exclude_recent;  /* NUM_EXCLUDE_RECENT 30 (used by the host mirror's tree bookkeeping) */
  int num_candidates;      /* NUM_CANDIDATES_FROM_TREE 3 (<= 16) */
  int tree_making_period;  /* TREE_MAKING_PERIOD_ 30 (host mirror) */
  double dist_thres;       /* SC_DIST_THRES 0.2 */
} ndtb_sc_params;
void ndtb_sc_default_params(ndtb_sc_params* p);

int ndtb_sc_create(int device, const ndtb_sc_params* p /* NULL: defaults */, ndtb_sc** out);
int ndtb_sc_destroy(ndtb_sc* s);
const char* ndtb_sc_last_error(const ndtb_sc* s);
int ndtb_sc_set_dist_thres(ndtb_sc* s, double thres);   /* SCManager::setSCdistThres (SC:262-265) */

/* n_scans clouds (pts[i]: n_points[i] records of stride_bytes, x y z float at offset 0; host memory) -> descriptors
 * first_id .. first_id + n_scans - 1, one launch for all of them. */
int ndtb_sc_add_scans(ndtb_sc* s, int n_scans, const void* const* pts, const size_t* n_points, size_t stride_bytes, int* first_id);
int ndtb_sc_size(const ndtb_sc* s);
int ndtb_sc_clear(ndtb_sc* s);
/* Descriptor id to the host: desc = num_ring x num_sector doubles COLUMN-major (Eigen's MatrixXd layout), ring key (num_ring
 * doubles), sector key (num_sector doubles), the float ring key the search uses (num_ring).  Any pointer may be NULL. */
int ndtb_sc_get(ndtb_sc* s, int id, double* desc, double* ring, double* sector, float* invkey);
/* distanceBtnScanContext(descriptor id1[i], descriptor id2[i]) -> dist[i], argmin column shift[i]; one block per pair. */
int ndtb_sc_distance(ndtb_sc* s, int n_pairs, const int* id1, const int* id2, double* dist, int* shift);
/* For each query q: candidates = the num_candidates nearest ring keys among descriptors [0, tree_size[q]), the one with the
 * smallest column-shift distance -> nn_idx[q], min_dist[q], yaw_diff_rad[q] = deg2rad(shift * 360 / num_sector) in float,
 * loop_id[q] = nn_idx[q] if min_dist[q] < dist_thres else -1.  tree_size[q] <= 0: loop_id -1, yaw 0 (SC:343-347). */
int ndtb_sc_detect(ndtb_sc* s, int n_queries, const int* query_id, const int* tree_size, int* loop_id, float* yaw_diff_rad,
                   double* min_dist, int* nn_idx);

#ifdef __cplusplus
}
#endif
#endif /* NDT_B200_SC_C_API_H_ */
