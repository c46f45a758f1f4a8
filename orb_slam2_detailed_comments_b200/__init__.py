"""B200-native ORB front-end: ORBextractor::operator() and ORBmatcher::DescriptorDistance k=2 scan of
cheukwaylee/ORB_SLAM2_detailed_comments, as hand-written sm_100a CUDA behind a C ABI (include/orb_b200.h).

Host-side mirrors of the reference classes (same names / argument meaning):
  ORBextractor  <- include/ORBextractor.h:80-250
  ORBmatcher    <- include/ORBmatcher.h:40-230 (DescriptorDistance + best/second-best/ratio semantics)
There is no CPU fallback; the CUDA library must be built (`__graft_entry__.build()`).
"""
from ._lib import KNN_DTYPE, KP_DTYPE, OrbError, build, lib  # noqa: F401
from .extractor import ORBextractor  # noqa: F401
from .matcher import ORBmatcher, hamming_knn, hamming_knn_sharded  # noqa: F401
from .vocabulary import ORBVocabulary  # noqa: F401
from .frontend import FrameFrontEnd  # noqa: F401
