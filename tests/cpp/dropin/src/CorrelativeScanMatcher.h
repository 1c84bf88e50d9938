// The reference's tests include "../src/CorrelativeScanMatcher.h": in this tree
// that path is the drop-in header.
#include "../../../../nautilus_b200/cpp/CorrelativeScanMatcher.h"
