"""nautilus_b200: B200 (sm_100a) correlative scan matching behind the reference's
CorrelativeScanMatcher interface.  The product is the C-ABI library
nautilus_b200/lib/libnautilus_csm_b200.so (include/nautilus_csm.h); the Python
modules are the host-side mirror used by tests and bench.py."""
from .matcher import CorrelativeScanMatcher, LookupTable, CsmError  # noqa: F401
from .point_matching import (PointCorrespondences, GetPointToPointMatching,  # noqa: F401
                             pose_transform, correspondences_from_matches)
