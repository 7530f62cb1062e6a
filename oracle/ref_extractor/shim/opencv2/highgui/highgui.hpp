// TEST INFRASTRUCTURE (oracle/ref_extractor): everything is in the core shim
#include <opencv2/core/core.hpp>
