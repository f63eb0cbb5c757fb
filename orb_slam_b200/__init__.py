"""orb_slam_b200 — B200-native ORB feature front end (drop-in for the
ORBextractor / ORBmatcher hot path of ORB-SLAM2).  Product code only; the CPU
oracle lives under /oracle and is never imported from here."""
from ._lib import KEYPOINT_DTYPE, FrontendError, LIB_PATH, lib  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
