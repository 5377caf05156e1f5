// stands in for <opencv2/imgproc/imgproc.hpp> when the reference extractor is compiled by oracle/ref_ext/make_ref_ext.py (test infrastructure)
#include "cv_shim.h"
