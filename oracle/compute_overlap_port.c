/* CPU restatement (plain C) of the reference's Cython IoU routine — TEST INFRASTRUCTURE ONLY.
 *
 * Follows /root/reference/PyraPose/utils/compute_overlap.pyx:13-53 operation for operation:
 * float64, the "+1" pixel convention, k outer / n inner, zero where the boxes do not overlap.
 * Built with -ffp-contract=off so no FMA is formed (the reference's x86-64 build has none).
 * Pinned against the reference extension itself (oracle/_ref) by tests/test_oracle_overlap.py.
 */
#include <stddef.h>

static inline double dmin(double a, double b) { return a < b ? a : b; }
static inline double dmax(double a, double b) { return a > b ? a : b; }

/* boxes: [n,4], query: [k,4], overlaps: [n,k] (row-major, fully written). */
void oracle_compute_overlap(const double* boxes, const double* query, double* overlaps, size_t n, size_t k) {
  for (size_t i = 0; i < n * k; ++i) overlaps[i] = 0.0;
  for (size_t q = 0; q < k; ++q) {
    const double* qb = query + 4 * q;
    const double box_area = (qb[2] - qb[0] + 1) * (qb[3] - qb[1] + 1);  /* pyx:32-35 */
    for (size_t i = 0; i < n; ++i) {
      const double* b = boxes + 4 * i;
      const double iw = dmin(b[2], qb[2]) - dmax(b[0], qb[0]) + 1;      /* pyx:37-40 */
      if (iw > 0) {
        const double ih = dmin(b[3], qb[3]) - dmax(b[1], qb[1]) + 1;    /* pyx:42-45 */
        if (ih > 0) {
          const double ua = (b[2] - b[0] + 1) * (b[3] - b[1] + 1) + box_area - iw * ih; /* pyx:47-51 */
          overlaps[i * k + q] = iw * ih / ua;                           /* pyx:52 */
        }
      }
    }
  }
}

/* Fused form of compute_gt_annotations (/root/reference/PyraPose/utils/anchors.py:310-316):
 * argmax over k (first maximum wins, as numpy.argmax), positive = max >= pos_thr,
 * ignore = max > neg_thr && !positive. */
void oracle_gt_assign(const double* boxes, const double* query, size_t n, size_t k, double neg_thr, double pos_thr,
                      long long* argmax, unsigned char* positive, unsigned char* ignore, double* max_overlap) {
  for (size_t i = 0; i < n; ++i) {
    const double* b = boxes + 4 * i;
    double best = -1.0;
    long long best_k = 0;
    for (size_t q = 0; q < k; ++q) {
      const double* qb = query + 4 * q;
      const double box_area = (qb[2] - qb[0] + 1) * (qb[3] - qb[1] + 1);
      double ov = 0.0;
      const double iw = dmin(b[2], qb[2]) - dmax(b[0], qb[0]) + 1;
      if (iw > 0) {
        const double ih = dmin(b[3], qb[3]) - dmax(b[1], qb[1]) + 1;
        if (ih > 0) {
          const double ua = (b[2] - b[0] + 1) * (b[3] - b[1] + 1) + box_area - iw * ih;
          ov = iw * ih / ua;
        }
      }
      if (ov > best) {
        best = ov;
        best_k = (long long)q;
      }
    }
    argmax[i] = best_k;
    if (max_overlap) max_overlap[i] = best;
    positive[i] = best >= pos_thr;
    ignore[i] = (best > neg_thr) && !(best >= pos_thr);
  }
}
