"""cv2 4.13 golden vectors for the CV_32F matrix algebra the reference's matcher uses (src/ORBmatcher.cc:1439 `Rcw*x3Dw+tcw`,
:1409 `-Rcw.t()*tcw`, :1189 `-sR21*t12`, :346 `cv::norm(PO)`), which oracle/ref_build/cvstub, the oracle and the product's
host code write out by hand.  Run HERE (needs the cv2 wheel):  python tests/golden/make_cv_algebra_golden.py"""
import os

import cv2
import numpy as np

assert cv2.__version__.startswith("4.13"), cv2.__version__
rng = np.random.default_rng(2024)
n = 4000
A = rng.normal(0, 1, (n, 3, 3)).astype(np.float32)
A[: n // 2] = np.linalg.qr(A[: n // 2].astype(np.float64))[0].astype(np.float32)          # half of them rotations
x = (rng.normal(0, 1, (n, 3, 1)) * rng.choice([0.01, 1, 30, 500], (n, 1, 1))).astype(np.float32)
c = (rng.normal(0, 1, (n, 3, 1)) * rng.choice([0.01, 1, 30], (n, 1, 1))).astype(np.float32)
out = np.zeros((n, 13), np.float32)
for i in range(n):
    out[i, 0:3] = cv2.gemm(A[i], x[i], 1.0, c[i], 1.0).ravel()                             # A*x + c   (MatExpr: one gemm, beta = 1)
    out[i, 3:6] = cv2.gemm(A[i], x[i], 1.0, None, 0.0).ravel()                             # A*x
    out[i, 6:9] = cv2.gemm(A[i], x[i], -1.0, None, 0.0, flags=cv2.GEMM_1_T).ravel()        # -A.t()*x  (MatExpr: alpha = -1, GEMM_1_T)
    out[i, 9:12] = cv2.gemm(A[i], x[i], -1.0, None, 0.0).ravel()                           # (-A)*x    (MatExpr: alpha = -1)
    out[i, 12] = np.float32(cv2.norm(x[i]))                                                # const float dist = cv::norm(PO)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cv_algebra.npz"), A=A, x=x, c=c, out=out)
print("wrote", n, "cases")
