"""Pose from votes — TEST INFRASTRUCTURE ONLY. The reference's own call, verbatim, on OpenCV (a third-party dependency
of the reference that IS installed here: cv2 4.13): /root/reference/PyraPose/utils/linemod_eval.py:427-430, 479-485."""
import numpy as np


def solve_pnp_ransac_votes(pose_votes, ori_points, camera, iterations=300, reprojection_error=5.0):
    """pose_votes (k, 16) float32; ori_points (8, 3) box corners; camera (fx, fy, cx, cy).
    Returns (retval, R (3,3), t (3,), inliers index array or None)."""
    import cv2
    k_hyp = pose_votes.shape[0]
    fx, fy, cx, cy = [float(v) for v in camera]
    K = np.float32([fx, 0., cx, 0., fy, cy, 0., 0., 1.]).reshape(3, 3)
    est_points = np.ascontiguousarray(pose_votes, dtype=np.float32).reshape((int(k_hyp * 8), 1, 2))
    obj_points = np.repeat(np.ascontiguousarray(ori_points, dtype=np.float32)[np.newaxis, :, :], k_hyp, axis=0)
    obj_points = obj_points.reshape((int(k_hyp * 8), 1, 3))
    retval, orvec, otvec, inliers = cv2.solvePnPRansac(objectPoints=obj_points, imagePoints=est_points, cameraMatrix=K,
                                                       distCoeffs=None, rvec=None, tvec=None, useExtrinsicGuess=False,
                                                       iterationsCount=iterations,
                                                       reprojectionError=reprojection_error, confidence=0.99,
                                                       flags=cv2.SOLVEPNP_ITERATIVE)
    if not retval:
        return False, None, None, None
    R_est, _ = cv2.Rodrigues(orvec)
    return True, R_est, otvec.reshape(3), inliers
