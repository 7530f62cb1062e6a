// TEST INFRASTRUCTURE (oracle/ref_matcher): include/Frame.h of the reference includes <opencv2/opencv.hpp>.
#include <opencv2/core/core.hpp>
