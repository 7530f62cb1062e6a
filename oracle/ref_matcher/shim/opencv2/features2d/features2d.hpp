// TEST INFRASTRUCTURE (oracle/ref_matcher): the reference includes this header for cv::KeyPoint, which the core shim defines.
#include <opencv2/core/core.hpp>
