// TEST INFRASTRUCTURE ONLY -- see opencv2/core/core.hpp in this directory.
#include "opencv2/core/core.hpp"
